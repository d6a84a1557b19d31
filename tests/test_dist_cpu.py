"""world_size-2 gloo tests (CPU) of the table-wise model-parallel plumbing in cloudtik_b200.ext_dist:
partition helpers, pooled-embedding all-to-all forward and its mirrored backward, uneven table split
(5 tables over 2 ranks) -- checked against the single-process global-batch computation, which is the
multi-GPU parity oracle."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _global_problem(n_tables, N, E):
    g = torch.Generator().manual_seed(1234)
    tables = [torch.randn(11 + 3 * t, E, generator=g) for t in range(n_tables)]
    idx = [torch.randint(0, 11 + 3 * t, (N,), generator=g) for t in range(n_tables)]
    gout = torch.randn(N, n_tables * E, generator=g)  # gradient w.r.t. the per-sample concatenated features
    return tables, idx, gout


def _worker(rank, world, port, n_tables, N, E, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from cloudtik_b200 import ext_dist
    from oracle import ops_numpy as ref
    ext_dist.init_distributed(rank=rank, size=world, backend="gloo")
    try:
        assert ext_dist.my_size == world and ext_dist.my_rank == rank
        sl = ext_dist.get_my_slice(n_tables)
        want_sl = ref.get_my_slice(n_tables, world, rank)
        assert (sl.start, sl.stop) == (want_sl.start, want_sl.stop)
        my_len, splits = ext_dist.get_split_lengths(n_tables)
        assert (my_len, splits) == ref.get_split_lengths(n_tables, world, rank)

        tables, idx, gout = _global_problem(n_tables, N, E)
        local = list(range(n_tables))[sl]
        lN = N // world
        # local tables looked up for the GLOBAL batch (one-hot), written as adjacent column blocks of
        # one [N, lS*E] buffer -- the layout the merged embedding kernel produces for the a2a
        buf = torch.empty(N, len(local) * E)
        for k, t in enumerate(local):
            buf[:, k * E:(k + 1) * E] = tables[t][idx[t]]
        buf.requires_grad_(True)
        ly = [buf[:, k * E:(k + 1) * E] for k in range(len(local))]
        req = ext_dist.alltoall(ly, splits)
        outs = req.wait()
        assert len(outs) == world
        got = torch.cat([o for o in outs], dim=1)            # [lN, n_tables*E], tables in global order
        want = torch.cat([tables[t][idx[t]] for t in range(n_tables)], dim=1)[rank * lN:(rank + 1) * lN]
        assert torch.equal(got, want)
        # backward: each rank feeds the gradient of ITS samples; the mirrored a2a must hand every rank the
        # gradient of its tables for the global batch
        got.backward(gout[rank * lN:(rank + 1) * lN])
        gw = torch.cat([gout[:, t * E:(t + 1) * E] for t in local], dim=1)
        assert torch.equal(buf.grad, gw)
        # shuffle_data: local-batch indices of all tables -> global-batch indices of my tables
        if n_tables % world == 0:
            loc_idx = [i[rank * lN:(rank + 1) * lN] for i in idx]
            sh = ext_dist.shuffle_data(loc_idx)
            g_o, g_i = ext_dist.global_batch_indices(sh, len(local))
            for k, t in enumerate(local):
                assert torch.equal(g_i[k], idx[t]) and torch.equal(g_o[k], torch.arange(N))
        # shuffle_bags: the generalised index all-to-all (uneven table split, ragged multi-hot bags, lengths travel)
        g = torch.Generator().manual_seed(77)
        bags = []
        for t in range(n_tables):
            lens = torch.randint(0, 4, (N,), generator=g)
            bags.append((torch.randint(0, 11 + 3 * t, (int(lens.sum()),), generator=g), torch.cumsum(lens, 0) - lens, lens))
        lo, li = [], []
        for ind, off, lens in bags:  # this rank's local slice of every table's bags
            a = int(off[rank * lN]) if lN else 0
            b = int(off[(rank + 1) * lN]) if (rank + 1) * lN < N else ind.numel()
            lo.append(off[rank * lN:(rank + 1) * lN] - a)
            li.append(ind[a:b])
        g_o, g_i = ext_dist.shuffle_bags(lo, li, splits)
        assert len(g_i) == len(local)
        for k, t in enumerate(local):
            assert torch.equal(g_i[k], bags[t][0]) and torch.equal(g_o[k], bags[t][1]), f"table {t}"
        # ... and its no-sync flavour for a fixed pooling factor
        P = 2
        fixed = [torch.randint(0, 11 + 3 * t, (N * P,), generator=g) for t in range(n_tables)]
        g_o, g_i = ext_dist.shuffle_bags([torch.arange(lN) * P] * n_tables,
                                         [f[rank * lN * P:(rank + 1) * lN * P] for f in fixed], splits, fixed_pooling=P)
        for k, t in enumerate(local):
            assert torch.equal(g_i[k], fixed[t]) and torch.equal(g_o[k], torch.arange(N) * P)
        # flat all-reduce with an early bucket reduced from inside backward (post-accumulate hooks)
        lin1, lin2 = torch.nn.Linear(3, 2), torch.nn.Linear(2, 1)
        with torch.no_grad():
            for p_ in list(lin1.parameters()) + list(lin2.parameters()):
                p_.fill_(0.5)
        allp = list(lin1.parameters()) + list(lin2.parameters())
        red = ext_dist.DenseGradAllreduce(allp, average=True, early=list(lin2.parameters()))
        for passes in (1, 2):
            for p_ in allp:
                p_.grad = None
            red.begin(passes)
            xs = [torch.full((4, 3), float(rank + 1 + q)) for q in range(passes)]
            for xq in xs:
                (lin2(lin1(xq)).sum() / passes).backward()
            red()
            # reference: mean over ranks of the same computation
            want = []
            for r in range(world):
                l1, l2 = torch.nn.Linear(3, 2), torch.nn.Linear(2, 1)
                with torch.no_grad():
                    for p_ in list(l1.parameters()) + list(l2.parameters()):
                        p_.fill_(0.5)
                for q in range(passes):
                    (l2(l1(torch.full((4, 3), float(r + 1 + q)))).sum() / passes).backward()
                want.append([p_.grad.clone() for p_ in list(l2.parameters()) + list(l1.parameters())])
            for k, p_ in enumerate(red.params):
                w = sum(wr[k] for wr in want) / world
                assert torch.allclose(p_.grad, w, rtol=1e-6, atol=1e-6), (passes, k)
        # flat all-reduce of the data-parallel MLP gradients (stand-in for DDP): mean and sum flavours
        ps = [torch.nn.Parameter(torch.zeros(3, 2)), torch.nn.Parameter(torch.zeros(5))]
        for average in (True, False):
            for k, p_ in enumerate(ps):
                p_.grad = torch.full_like(p_, float((rank + 1) * (k + 1)))
            ext_dist.DenseGradAllreduce(ps, average=average)()
            tot = sum(r + 1 for r in range(world))
            for k, p_ in enumerate(ps):
                want = tot * (k + 1) / (world if average else 1)
                assert torch.equal(p_.grad, torch.full_like(p_, want))
        # all_gather of per-rank scores (eval path)
        scores = torch.full((lN, 1), float(rank))
        allg = ext_dist.all_gather(scores, None)
        assert torch.equal(allg, torch.arange(world, dtype=torch.float32).repeat_interleave(lN).view(-1, 1))
        ext_dist.barrier()
        ret[rank] = "ok"
    finally:
        torch.distributed.destroy_process_group()


@pytest.mark.parametrize("n_tables,N", [(5, 6), (4, 8)])
def test_alltoall_two_ranks_gloo(n_tables, N):
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_tables, N, 3, ret), nprocs=world, join=True)
    assert dict(ret) == {0: "ok", 1: "ok"}


def test_rowwise_filter_owned_matches_per_bag_filter():
    """cloudtik_b200.rowwise.filter_owned (pure index arithmetic, runs on CPU): bags over global row ids ->
    bags over the rows rank r owns (id % W == r, local id = id // W), empty bags preserved."""
    from cloudtik_b200.rowwise import filter_owned
    rs = np.random.RandomState(0)
    for W in (2, 3, 8):
        lens = rs.randint(0, 6, size=40)
        off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        idx = rs.randint(0, 1000, size=int(lens.sum())).astype(np.int64)
        for r in range(W):
            li, lo = filter_owned(torch.tensor(idx), torch.tensor(off), W, r)
            want_i, want_o = [], [0]
            for b in range(len(lens)):
                bag = idx[off[b]:off[b + 1]]
                keep = bag[bag % W == r] // W
                want_i += keep.tolist()
                want_o.append(len(want_i))
            assert li.tolist() == want_i and lo.tolist() == want_o
