"""Host-side logic of the micro-batched multi-GPU step (no GPU needed): the cut of a global batch into
micro-batches keeps every bag, in the order the sharded kernels expect."""
import numpy as np
import torch

from cloudtik_b200.p2p import split_global_batch


def _bags(rs, B, rows, max_len):
    lens = rs.randint(0, max_len + 1, size=B)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    idx = rs.randint(0, rows, size=int(lens.sum())).astype(np.int64)
    return torch.tensor(idx), torch.tensor(off), lens


def test_split_global_batch_ragged():
    rs = np.random.RandomState(0)
    world, n_chunks, lN = 4, 2, 6
    B = world * lN
    idx_l, off_l, lens_l = [], [], []
    for rows in (50, 7, 1000):
        i, o, l = _bags(rs, B, rows, 4)
        idx_l.append(i), off_l.append(o), lens_l.append(l)
    chunks = split_global_batch(idx_l, off_l, world, n_chunks)
    assert len(chunks) == n_chunks
    per = lN // n_chunks
    for c, (ic, oc) in enumerate(chunks):
        for t in range(3):
            # bags of chunk c, rank-major: rank r contributes its samples [c*per, (c+1)*per)
            want_bags = [r * lN + c * per + k for r in range(world) for k in range(per)]
            assert oc[t].numel() == len(want_bags)
            ends = np.concatenate([off_l[t].numpy()[1:], [idx_l[t].numel()]])
            pos = 0
            for j, b in enumerate(want_bags):
                s, e = int(off_l[t][b]), int(ends[b])
                assert int(oc[t][j]) == pos
                np.testing.assert_array_equal(ic[t][pos:pos + (e - s)].numpy(), idx_l[t][s:e].numpy())
                pos += e - s
            assert pos == ic[t].numel()


def test_split_global_batch_one_chunk_is_identity():
    rs = np.random.RandomState(1)
    i, o, _ = _bags(rs, 8, 20, 3)
    (ic, oc), = split_global_batch([i], [o], 2, 1)
    assert torch.equal(ic[0], i) and torch.equal(oc[0], o)
