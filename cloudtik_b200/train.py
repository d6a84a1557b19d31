"""DLRM training entry point on the B200 kernels: ``python -m cloudtik_b200.train <reference flags>``.

Drop-in for the way CloudTik's quick-start launches the reference
(``QS/scripts/recommendation/pytorch/dlrm/training/cpu/training.sh:83-93`` ->
``cloudtik-run ... dlrm_s_pytorch.py --arch-... --mini-batch-size ... --learning-rate ...``): the flag names and
DEFAULTS of the training variant (``T/dlrm_s_pytorch.py:640-742``) for everything on the hot path -- model shape,
``--arch-interaction-op`` / ``--arch-interaction-itself``, ``--loss-function mse|bce|wbce``, the random data
generator's flags, LR policy, ``--sparse-dense-boundary`` -- the same progress line
(``Finished training it j/N of epoch k, X ms/it, loss L, accuracy A %``, ``T/dlrm_s_pytorch.py:1277-1283``) and the
final ``Throughput: X fps`` line the shell wrapper of the product variant greps (``P/dlrm_s_pytorch.py:964-970``).

Data: ``--data-generation=random|synthetic`` (this image has no Criteo files), through the mirror of the reference's
``RandomDataset`` (``cloudtik_b200.data``): the same numpy stream in the same order, and the tables drawn from the
per-table seeds like ``T/dlrm_s_pytorch.py:240-243``.  With the reference's own test flags
(``T/test/dlrm_s_test.sh:24``: ``--mini-batch-size=2 --data-size=4 --nepochs=1 --arch-interaction-op=dot
--learning-rate=0.1``) this prints the reference's losses, 0.028790 and 0.156787.

Multi-GPU: launch one process per GPU (torchrun / ``python -m cloudtik_b200.launcher``); tables at or above
``--sparse-dense-boundary`` rows are sharded table-wise, smaller ones replicated, every rank is fed its local batch
of every table and the sharded tables' bags travel through the index all-to-all, as in the reference's
``distributed_forward`` (``T/dlrm_s_pytorch.py:483-522``).
"""
from __future__ import annotations

import argparse
import sys
import time

import numpy as np
import torch

from . import data as rnd_data
from . import ext_dist
from .dlrm import DLRM_Net, enable_peer_exchange
from .lr import LRPolicyScheduler


def dash_separated_ints(value):
    for v in value.split("-"):
        int(v)
    return value


def parse(argv=None):
    p = argparse.ArgumentParser(description="Train Deep Learning Recommendation Model (DLRM) on B200")
    # model (T/dlrm_s_pytorch.py:646-655)
    p.add_argument("--arch-sparse-feature-size", type=int, default=2)
    p.add_argument("--arch-embedding-size", type=dash_separated_ints, default="4-3-2")
    p.add_argument("--arch-mlp-bot", type=dash_separated_ints, default="4-3-2")
    p.add_argument("--arch-mlp-top", type=dash_separated_ints, default="4-2-1")
    p.add_argument("--arch-interaction-op", type=str, default="dot")
    p.add_argument("--arch-interaction-itself", action="store_true", default=False)
    # loss (T/dlrm_s_pytorch.py:671-675)
    p.add_argument("--loss-function", type=str, default="mse")  # or bce or wbce
    p.add_argument("--loss-weights", type=str, default="1.0-1.0")  # for wbce
    p.add_argument("--loss-threshold", type=float, default=0.0)
    p.add_argument("--round-targets", type=bool, default=False)
    # data (random generator flags, T/dlrm_s_pytorch.py:677-697)
    p.add_argument("--data-size", type=int, default=1)
    p.add_argument("--num-batches", type=int, default=0)
    p.add_argument("--data-generation", type=str, default="random")  # random | synthetic (trace-driven)
    p.add_argument("--data-trace-file", type=str, default="./input/dist_emb_j.log")
    p.add_argument("--data-trace-enable-padding", type=bool, default=False)
    p.add_argument("--num-indices-per-lookup", type=int, default=10)
    p.add_argument("--num-indices-per-lookup-fixed", type=bool, default=False)
    # training (T/dlrm_s_pytorch.py:701-707)
    p.add_argument("--mini-batch-size", type=int, default=1)
    p.add_argument("--nepochs", type=int, default=1)
    p.add_argument("--learning-rate", type=float, default=0.01)
    p.add_argument("--print-precision", type=int, default=5)
    p.add_argument("--numpy-rand-seed", type=int, default=123)
    p.add_argument("--print-freq", type=int, default=1)
    p.add_argument("--print-time", action="store_true", default=False)
    p.add_argument("--debug-mode", action="store_true", default=False)
    p.add_argument("--num-warmup-iters", type=int, default=0)
    p.add_argument("--lr-num-warmup-steps", type=int, default=0)
    p.add_argument("--lr-decay-start-step", type=int, default=0)
    p.add_argument("--lr-num-decay-steps", type=int, default=0)
    p.add_argument("--sparse-dense-boundary", type=int, default=2048)
    p.add_argument("--bf16", action="store_true", default=False)
    p.add_argument("--save-model", type=str, default="")
    p.add_argument("--load-model", type=str, default="")
    p.add_argument("--dist-backend", "--backend", dest="backend", type=str, default="nccl")
    # beyond the reference
    p.add_argument("--embedding-optimizer", type=str, default="sgd", choices=["sgd", "adagrad"])
    p.add_argument("--no-p2p", action="store_true", help="multi-GPU: NCCL all-to-all instead of peer memory")
    return p.parse_args(argv)


def _ckpt_path(path, rank, world):
    """Rank 0 (and a single-process run) uses the given name, the other ranks append ``.rank<r>``."""
    return path if world == 1 or rank == 0 else "{}.rank{}".format(path, rank)


def make_loss(args, dev):
    """``loss_fn_wrap`` of the training variant (T/dlrm_s_pytorch.py:1000-1008,1058-1075)."""
    if args.loss_function == "mse":
        fn = torch.nn.MSELoss(reduction="mean")
        return lambda Z, T: fn(Z, T)
    if args.loss_function == "bce":
        fn = torch.nn.BCELoss(reduction="mean")
        return lambda Z, T: fn(Z, T)
    if args.loss_function == "wbce":
        ws = torch.tensor(np.fromstring(args.loss_weights, dtype=float, sep="-"), device=dev)
        fn = torch.nn.BCELoss(reduction="none")
        return lambda Z, T: (ws[T.detach().view(-1).long()].view_as(T) * fn(Z, T)).mean()
    sys.exit("ERROR: --loss-function=" + args.loss_function + " is not supported")


def run(argv=None):
    args = parse(argv)
    if args.data_generation not in ("random", "synthetic"):
        sys.exit("ERROR: --data-generation=" + args.data_generation + " is not available (no dataset files in this build)")
    ext_dist.init_distributed(backend=args.backend)
    world, rank = max(ext_dist.my_size, 1), max(ext_dist.my_rank, 0)
    dev = torch.device("cuda", torch.cuda.current_device())
    np.random.seed(args.numpy_rand_seed)  # T/dlrm_s_pytorch.py:743-746
    np.set_printoptions(precision=args.print_precision)
    torch.set_printoptions(precision=args.print_precision)
    torch.manual_seed(args.numpy_rand_seed)

    ln_bot = np.fromstring(args.arch_mlp_bot, dtype=int, sep="-")
    ln_emb = np.fromstring(args.arch_embedding_size, dtype=int, sep="-")
    m_den = ln_bot[0]
    # the data loader comes first, as in the reference (T/dlrm_s_pytorch.py:811); it draws nothing until read
    train_data, train_ld = rnd_data.make_random_data_and_loader(args, ln_emb, m_den, world_size=world)
    nbatches = args.num_batches if args.num_batches > 0 else len(train_ld)

    m_spa = args.arch_sparse_feature_size
    num_fea = ln_emb.size + 1
    m_den_out = ln_bot[ln_bot.size - 1]
    if args.arch_interaction_op == "dot":  # T/dlrm_s_pytorch.py:873-881
        num_int = (num_fea * (num_fea + 1)) // 2 + m_den_out if args.arch_interaction_itself else \
            (num_fea * (num_fea - 1)) // 2 + m_den_out
    elif args.arch_interaction_op == "cat":
        num_int = num_fea * m_den_out
    else:
        sys.exit("ERROR: --arch-interaction-op=" + args.arch_interaction_op + " is not supported")
    ln_top = np.fromstring(str(num_int) + "-" + args.arch_mlp_top, dtype=int, sep="-")
    if m_den != ln_bot[0] or m_spa != m_den_out:
        sys.exit("ERROR: arch-sparse-feature-size %d does not match last dim of bottom mlp %d" % (m_spa, m_den_out))

    use_p2p = world > 1 and not args.no_p2p
    dlrm = DLRM_Net(m_spa, ln_emb, ln_bot, ln_top, sigmoid_bot=-1, sigmoid_top=ln_top.size - 2,
                    loss_threshold=args.loss_threshold, np_init_emb_weight=ln_emb.sum() * m_spa < (1 << 28),
                    np_emb_seeds=True, device=dev, lr=args.learning_rate / world,
                    emb_optimizer=args.embedding_optimizer, bf16=args.bf16,
                    arch_interaction_op=args.arch_interaction_op, arch_interaction_itself=args.arch_interaction_itself,
                    sparse_dense_boundary=args.sparse_dense_boundary if use_p2p else 0)
    if args.debug_mode and rank == 0:
        print("model arch:")
        print(dlrm)
    B = args.mini_batch_size  # global batch; RandomDataset hands every rank B // world samples
    lB = B // world
    local = list(range(ln_emb.size)) if world == 1 else dlrm.local_ln_emb
    if use_p2p:
        enable_peer_exchange(dlrm, lB)
    loss_fn = make_loss(args, dev)
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

    total_time = total_loss = total_accu = 0.0
    total_iter = total_samp = 0
    perf_time, perf_iters = 0.0, 0
    for k in range(args.nepochs):
        for j, (X, lS_o, lS_i, T) in enumerate(train_ld):
            if j < start_iter:
                continue  # the stream is replayed up to the checkpointed iteration, so a resumed run sees the same data
            if j >= nbatches:
                break
            X, T = X.to(dev), T.to(dev)
            lS_o = [o.to(dev) for o in lS_o]
            lS_i = [i.to(dev) for i in lS_i]
            torch.cuda.synchronize()
            t1 = time.time()
            # lr is re-read every step, as the fused path of the reference does (P/dlrm_s_pytorch.py:1027)
            dlrm.emb_l.sgd_args = dlrm.emb_l.sgd_args._replace(lr=lr_scheduler.get_last_lr()[0] / world)
            Z = dlrm(X, lS_o, lS_i).float()
            E = loss_fn(Z, T)
            L = E.detach().cpu().numpy()  # per-step loss read, like the reference (T/dlrm_s_pytorch.py:1223)
            S = Z.detach().cpu().numpy()
            optimizer.zero_grad(set_to_none=True)
            E.backward()
            sync_dense()
            optimizer.step()
            lr_scheduler.step()
            torch.cuda.synchronize()
            t2 = time.time()
            total_time += t2 - t1
            mbs = T.shape[0]
            total_accu += float(np.sum((np.round(S, 0) == T.cpu().numpy()).astype(np.uint8)))
            total_loss += float(L) * mbs
            total_iter += 1
            total_samp += mbs
            if ((j + 1) % args.print_freq == 0) or (j + 1 == nbatches):
                gT = 1000.0 * total_time / total_iter if args.print_time else -1
                if j + 1 > args.num_warmup_iters:
                    perf_time += total_time
                    perf_iters += total_iter
                if rank == 0:
                    print("Finished training it {}/{} of epoch {}, {:.2f} ms/it, ".format(j + 1, nbatches, k, gT)
                          + "loss {:.6f}, accuracy {:3.3f} %".format(total_loss / total_samp,
                                                                   100.0 * total_accu / total_samp), flush=True)
                total_time = total_loss = total_accu = 0.0
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
