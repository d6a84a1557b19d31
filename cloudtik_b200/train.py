"""DLRM training entry point on the B200 kernels: ``python -m cloudtik_b200.train <reference flags>``.

Drop-in for the way CloudTik's quick-start launches the reference
(``QS/scripts/recommendation/pytorch/dlrm/training/cpu/training.sh:83-93`` ->
``cloudtik-run ... P/dlrm_s_pytorch.py --arch-... --mini-batch-size ... --learning-rate ...``): the same flag
names for everything on the hot path, the same progress line
(``Finished training it j/N of epoch k, X ms/it, loss L``, ``P/dlrm_s_pytorch.py:1094-1101``) and the same
final ``Throughput: X fps`` line the shell wrapper greps (``P/dlrm_s_pytorch.py:964-970``).

Data: synthetic only (``--data-generation=random`` of ``T/dlrm_s_pytorch.py``): this image has no Criteo
files; the loaders are SURVEY.md section 8f "next".  Multi-GPU: launch one process per GPU (torchrun /
cloudtik-run env), tables are sharded table-wise exactly like the reference's ``distributed_forward``.
"""
from __future__ import annotations

import argparse
import sys
import time

import numpy as np
import torch

from . import ext_dist
from .dlrm import DLRM_Net, enable_peer_exchange
from .lr import LRPolicyScheduler


def dash_separated_ints(value):
    for v in value.split("-"):
        int(v)
    return value


def parse(argv=None):
    p = argparse.ArgumentParser(description="Train Deep Learning Recommendation Model (DLRM) on B200")
    # model (P/dlrm_s_pytorch.py:647-657)
    p.add_argument("--arch-sparse-feature-size", type=int, default=2)
    p.add_argument("--arch-embedding-size", type=dash_separated_ints, default="4-3-2")
    p.add_argument("--arch-mlp-bot", type=dash_separated_ints, default="4-3-2")
    p.add_argument("--arch-mlp-top", type=dash_separated_ints, default="4-2-1")
    p.add_argument("--loss-threshold", type=float, default=0.0)
    p.add_argument("--round-targets", type=bool, default=False)
    # data (random generator flags of T/dlrm_s_pytorch.py:668-676)
    p.add_argument("--data-generation", type=str, default="random")
    p.add_argument("--num-batches", type=int, default=0)
    p.add_argument("--num-indices-per-lookup", type=int, default=1)
    p.add_argument("--num-indices-per-lookup-fixed", type=bool, default=True)
    # training (P/dlrm_s_pytorch.py:679-684)
    p.add_argument("--mini-batch-size", type=int, default=1)
    p.add_argument("--nepochs", type=int, default=1)
    p.add_argument("--learning-rate", type=float, default=0.01)
    p.add_argument("--numpy-rand-seed", type=int, default=123)
    p.add_argument("--print-freq", type=int, default=1)
    p.add_argument("--print-time", action="store_true", default=False)
    p.add_argument("--num-warmup-iters", type=int, default=0)
    p.add_argument("--lr-num-warmup-steps", type=int, default=0)
    p.add_argument("--lr-decay-start-step", type=int, default=0)
    p.add_argument("--lr-num-decay-steps", type=int, default=0)
    p.add_argument("--bf16", action="store_true", default=False)
    p.add_argument("--save-model", type=str, default="")
    p.add_argument("--load-model", type=str, default="")
    p.add_argument("--backend", type=str, default="nccl")
    p.add_argument("--embedding-optimizer", type=str, default="sgd", choices=["sgd", "adagrad"])
    p.add_argument("--no-p2p", action="store_true", help="multi-GPU: NCCL all-to-all instead of peer memory")
    return p.parse_args(argv)


def random_batch(rs, B, ln_emb_local, m_den, P, fixed, round_targets, dev):
    """One synthetic batch: dense U[0,1), one bag of up to P uniform indices per (sample, table)."""
    X = torch.from_numpy(rs.rand(B, m_den).astype(np.float32)).to(dev)
    T = rs.rand(B, 1).astype(np.float32)
    T = torch.from_numpy(np.round(T) if round_targets else T).to(dev)
    idx_l, off_l = [], []
    for n in ln_emb_local:
        lens = np.full(B, P) if fixed else rs.randint(1, P + 1, size=B)
        off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
        idx = np.floor(rs.rand(int(lens.sum())) * n).astype(np.int64)
        idx_l.append(torch.from_numpy(idx).to(dev))
        off_l.append(torch.from_numpy(off).to(dev))
    return X, T, idx_l, off_l


def _ckpt_path(path, rank, world):
    """Rank 0 (and a single-process run) uses the given name, the other ranks append ``.rank<r>``."""
    return path if world == 1 or rank == 0 else "{}.rank{}".format(path, rank)


def run(argv=None):
    args = parse(argv)
    if args.data_generation != "random":
        sys.exit("ERROR: only --data-generation=random is available (no dataset files in this build)")
    ext_dist.init_distributed(backend=args.backend)
    world, rank = max(ext_dist.my_size, 1), max(ext_dist.my_rank, 0)
    dev = torch.device("cuda", torch.cuda.current_device())
    np.random.seed(args.numpy_rand_seed)
    torch.manual_seed(args.numpy_rand_seed)

    ln_bot = np.fromstring(args.arch_mlp_bot, dtype=int, sep="-")
    ln_emb = np.fromstring(args.arch_embedding_size, dtype=int, sep="-")
    m_spa = args.arch_sparse_feature_size
    num_fea = ln_emb.size + 1
    num_int = (num_fea * (num_fea - 1)) // 2 + m_spa  # P/dlrm_s_pytorch.py:779 (dot, no self interaction)
    ln_top = np.fromstring(str(num_int) + "-" + args.arch_mlp_top, dtype=int, sep="-")
    if m_spa != ln_bot[-1]:
        sys.exit("ERROR: arch-sparse-feature-size %d does not match last dim of bottom mlp %d" % (m_spa, ln_bot[-1]))

    dlrm = DLRM_Net(m_spa, ln_emb, ln_bot, ln_top, sigmoid_bot=-1, sigmoid_top=ln_top.size - 2,
                    loss_threshold=args.loss_threshold, np_init_emb_weight=ln_emb.sum() * m_spa < (1 << 28),
                    device=dev, lr=args.learning_rate / world, emb_optimizer=args.embedding_optimizer,
                    bf16=args.bf16)
    B = args.mini_batch_size  # global batch (the reference slices it by rank, P/dlrm_s_pytorch.py:990-997)
    lB = B // world
    local = list(range(ln_emb.size)) if world == 1 else dlrm.local_ln_emb
    if world > 1 and not args.no_p2p:
        enable_peer_exchange(dlrm, lB)
    dense = [p for p in dlrm.parameters() if p.requires_grad]
    optimizer = torch.optim.SGD(dense, lr=args.learning_rate)
    if args.bf16:
        from .mlp import enable_bf16_shadows
        enable_bf16_shadows([dlrm.bot_l, dlrm.top_l], optimizer)
    sync_dense = ext_dist.DenseGradAllreduce(dense) if world > 1 else (lambda: None)
    lr_scheduler = LRPolicyScheduler(optimizer, args.lr_num_warmup_steps, args.lr_decay_start_step,
                                     args.lr_num_decay_steps)
    start_iter = 0
    if args.load_model:
        # every rank reads its own shard (tables are table-wise sharded; the MLPs are replicated in each shard)
        ld = torch.load(_ckpt_path(args.load_model, rank, world), map_location=dev)
        dlrm.load_state_dict(ld["state_dict"])
        optimizer.load_state_dict(ld["opt_state_dict"])
        if ld.get("emb_opt_state") is not None:
            dlrm.emb_l.load_optimizer_state_dict(ld["emb_opt_state"])
        if ld.get("lr_sched_state") is not None:
            lr_scheduler.load_state_dict(ld["lr_sched_state"])
        start_iter = ld.get("iter", 0)
        if rank == 0:
            print("Loaded model {} (iter {})".format(args.load_model, start_iter))

    nbatches = args.num_batches if args.num_batches > 0 else 1
    rs = np.random.RandomState(args.numpy_rand_seed + 1)
    total_time = total_loss = 0.0
    total_iter = total_samp = 0
    perf_time, perf_iters = 0.0, 0
    for k in range(args.nepochs):
        for j in range(start_iter, nbatches):
            X, T, idx_l, off_l = random_batch(rs, B, ln_emb, ln_bot[0], args.num_indices_per_lookup,
                                              args.num_indices_per_lookup_fixed, args.round_targets, dev)
            if world > 1:
                X, T = X[rank * lB:(rank + 1) * lB], T[rank * lB:(rank + 1) * lB]
            indices, offsets, iwro = dlrm.emb_l.linearize_indices_and_offsets(
                [idx_l[t] for t in local], [off_l[t] for t in local], [False] * len(local))
            torch.cuda.synchronize()
            t1 = time.time()
            # lr is re-read every step, as the fused path of the reference does (P/dlrm_s_pytorch.py:1027)
            dlrm.emb_l.sgd_args = dlrm.emb_l.sgd_args._replace(lr=lr_scheduler.get_last_lr()[0] / world)
            Z = dlrm(X, indices, offsets, iwro).float()
            E = dlrm.loss_fn(Z, T)
            L = E.detach().cpu().numpy()  # per-step loss read, like the reference (P/dlrm_s_pytorch.py:1048)
            optimizer.zero_grad(set_to_none=True)
            E.backward()
            sync_dense()
            optimizer.step()
            lr_scheduler.step()
            torch.cuda.synchronize()
            t2 = time.time()
            total_time += t2 - t1
            mbs = T.shape[0]
            total_loss += float(L) * mbs
            total_iter += 1
            total_samp += mbs
            if ((j + 1) % args.print_freq == 0) or (j + 1 == nbatches):
                gT = 1000.0 * total_time / total_iter if args.print_time else -1
                if j + 1 > args.num_warmup_iters:
                    perf_time += total_time
                    perf_iters += total_iter
                if rank == 0:
                    print("Finished training it {}/{} of epoch {}, {:.2f} ms/it,".format(j + 1, nbatches, k, gT)
                          + " loss {:.6f}".format(total_loss / total_samp), flush=True)
                total_time = total_loss = 0.0
                total_iter = total_samp = 0
        start_iter = 0
    if rank == 0 and perf_iters > 0:
        print("Throughput: {:.3f} fps".format(perf_iters * B / perf_time))
    if args.save_model:
        # checkpoint schema of T/dlrm_s_pytorch.py:1408-1423 (state_dict keys emb_l.{k}.weight, bot_l..., top_l...)
        # plus what the fused update keeps outside the weights (Adagrad sums, bf16 trails) and the LR policy state.
        # One file per rank when the tables are sharded: "local_tables" names the global ids behind emb_l.0..k.
        torch.save({"epoch": args.nepochs, "iter": nbatches, "nepochs": args.nepochs, "nbatches": nbatches,
                    "state_dict": dlrm.state_dict(), "opt_state_dict": optimizer.state_dict(),
                    "emb_opt_state": dlrm.emb_l.optimizer_state_dict(),
                    "lr_sched_state": lr_scheduler.state_dict(), "world": world, "rank": rank,
                    "local_tables": [int(t) for t in local]}, _ckpt_path(args.save_model, rank, world))
        if rank == 0:
            print("Saving model to {}".format(args.save_model))
    if world > 1:
        ext_dist.barrier()
        torch.distributed.destroy_process_group()
    return dlrm


if __name__ == "__main__":
    run()
