"""Multi-GPU parity run (one process per GPU, NCCL): table-wise sharded DLRM steps vs the single-process
global-batch CPU oracle (that run IS the multi-GPU oracle).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tools/dist_parity.py [--bf16] [--tables 5]

Every rank builds the same oracle port on the CPU, copies its slice of the tables (and the MLPs) into the
B200 model, trains `--steps` steps on its slice of the global batch and checks: mean-of-rank losses vs the
oracle's global-batch loss, its own tables and the MLP weights vs the oracle's, fp32 <= 1e-5 / bf16 <= 1e-2.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bf16", action="store_true")
    ap.add_argument("--tables", type=int, default=8)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--batch", type=int, default=512, help="GLOBAL batch")
    ap.add_argument("--ddp", action="store_true", help="wrap the MLPs in DDP instead of the flat all-reduce")
    ap.add_argument("--p2p", action="store_true", help="fused embedding + all-to-all over NVLink peer memory")
    a = ap.parse_args()

    from cloudtik_b200 import ext_dist
    from cloudtik_b200.dlrm import DLRM_Net, enable_peer_exchange
    from oracle.dlrm_port import DLRMPort, train_step as cpu_step, uniform_batch

    ext_dist.init_distributed(backend="nccl")
    world, rank = ext_dist.my_size, ext_dist.my_rank
    dev = torch.device("cuda", torch.cuda.current_device())
    D = 128
    rows = [5000, 37, 20000, 3, 1200, 800, 64, 100000][:a.tables]
    F = len(rows) + 1
    ln_bot, ln_top = [13, 256, D], [D + F * (F - 1) // 2, 512, 256, 1]
    lr = 0.05
    np.random.seed(11)
    port = DLRMPort(D, rows, ln_bot, ln_top)
    np.random.seed(11)
    net = DLRM_Net(D, np.array(rows), np.array(ln_bot), np.array(ln_top), sigmoid_top=len(ln_top) - 2,
                   np_init_emb_weight=True, device=dev, lr=lr / world, bf16=a.bf16)
    local = net.local_ln_emb
    with torch.no_grad():
        for k, t in enumerate(local):
            net.emb_l.weights[k].copy_(port.emb_l[t].weight)
        for dst, src in ((net.bot_l, port.bot_l), (net.top_l, port.top_l)):
            for pd, ps in zip(dst.parameters(), src.parameters()):
                pd.copy_(ps)
    if a.p2p:
        enable_peer_exchange(net, a.batch // world)
    if a.ddp:
        net.bot_l = ext_dist.DDP(net.bot_l, device_ids=[dev.index])
        net.top_l = ext_dist.DDP(net.top_l, device_ids=[dev.index])
    copt = torch.optim.SGD(port.parameters(), lr=lr)
    dense = [p for p in net.parameters() if p.requires_grad]
    gopt = torch.optim.SGD(dense, lr=lr)
    sync_dense = (lambda: None) if a.ddp else ext_dist.DenseGradAllreduce(dense)
    rs = np.random.RandomState(5)
    N, lN = a.batch, a.batch // world
    tol = 1e-2 if a.bf16 else 1e-5
    for step in range(a.steps):
        X, lS_o, lS_i, T = uniform_batch(rs, N, rows, pooling=3, unique_in_bag=(step % 2 == 1))
        want = cpu_step(port, copt, torch.tensor(X), [torch.tensor(o) for o in lS_o],
                        [torch.tensor(i) for i in lS_i], torch.tensor(T), bf16=a.bf16)
        Xl = torch.tensor(X[rank * lN:(rank + 1) * lN]).to(dev)
        Tl = torch.tensor(T[rank * lN:(rank + 1) * lN]).to(dev)
        lo = [torch.tensor(lS_o[t]).to(dev) for t in local]
        li = [torch.tensor(lS_i[t]).to(dev) for t in local]
        Z = net(Xl, lo, li).float()
        E = net.loss_fn(Z, Tl)
        gopt.zero_grad(set_to_none=True)
        E.backward()
        sync_dense()
        gopt.step()
        Lm = E.detach().clone()
        torch.distributed.all_reduce(Lm)
        got = float(Lm.cpu()) / world
        assert abs(got - want) <= tol * abs(want), f"step {step}: loss {got} vs oracle {want}"
    for k, t in enumerate(local):
        w = net.emb_l.weights[k].float().cpu().numpy()
        np.testing.assert_allclose(w, port.emb_l[t].weight.detach().numpy(), rtol=tol,
                                   atol=2e-3 if a.bf16 else 1e-6, err_msg=f"table {t}")  # abs: rows near zero
    top = net.top_l.module if a.ddp else net.top_l
    for (n1, p1), p2 in zip(top.named_parameters(), port.top_l.parameters()):
        np.testing.assert_allclose(p1.detach().cpu().numpy(), p2.detach().numpy(), rtol=10 * tol,
                                   atol=2e-3 if a.bf16 else 1e-6, err_msg=n1)
    ext_dist.barrier()
    if rank == 0:
        print(f"dist parity ok: world={world} tables={len(rows)} bf16={a.bf16} p2p={a.p2p} steps={a.steps} "
              f"(losses and weights match the single-process oracle)")
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
