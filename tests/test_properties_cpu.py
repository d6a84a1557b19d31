"""Property tests (hypothesis) of the CPU oracle and of the pure host-side index arithmetic: the checker itself must
hold the invariants the GPU parity tests lean on."""
import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from oracle import ops_numpy as ref

_bags = st.lists(st.integers(min_value=0, max_value=6), min_size=1, max_size=40)


@settings(max_examples=60, deadline=None)
@given(lens=_bags, rows=st.integers(min_value=1, max_value=50), seed=st.integers(0, 2**31 - 1))
def test_pooling_is_linear_and_matches_offset2bag(lens, rows, seed):
    rs = np.random.RandomState(seed)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    idx = rs.randint(0, rows, size=int(sum(lens))).astype(np.int64)
    W1, W2 = rs.randn(rows, 4).astype(np.float32), rs.randn(rows, 4).astype(np.float32)
    a, b = ref.embedding_bag_sum(W1, idx, off), ref.embedding_bag_sum(W2, idx, off)
    np.testing.assert_allclose(ref.embedding_bag_sum(W1 + W2, idx, off), a + b, rtol=1e-5, atol=1e-5)
    # bag membership: scattering ones through offset2bag reproduces the bag lengths; empty bags pool to zero
    bag = ref.offset2bag(off, len(idx))
    np.testing.assert_array_equal(np.bincount(bag, minlength=len(lens)), np.array(lens))
    np.testing.assert_array_equal(ref.bag_lengths(off, len(idx)), np.array(lens))
    assert not a[np.array(lens) == 0].any()


@settings(max_examples=60, deadline=None)
@given(keys=st.lists(st.integers(min_value=0, max_value=200), min_size=0, max_size=300))
def test_sort_dedup_oracle_matches_numpy(keys):
    k = np.array(keys, dtype=np.uint32)
    sk, perm, uq, counts, seg = ref.sort_dedup(k)
    np.testing.assert_array_equal(sk, np.sort(k, kind="stable"))
    np.testing.assert_array_equal(k[perm], sk)
    if len(k):
        same = sk[1:] == sk[:-1]
        assert (perm[1:][same] > perm[:-1][same]).all()          # stable
    u, c = np.unique(k, return_counts=True)
    np.testing.assert_array_equal(uq, u)
    np.testing.assert_array_equal(np.diff(seg), c)
    np.testing.assert_array_equal(counts, c)


@settings(max_examples=100, deadline=None)
@given(n=st.integers(min_value=0, max_value=200), world=st.integers(min_value=1, max_value=16))
def test_table_partition_tiles_the_range(n, world):
    cover = []
    for r in range(world):
        sl = ref.get_my_slice(n, world, r)
        cover += list(range(n))[sl]
        my_len, splits = ref.get_split_lengths(n, world, r)
        assert my_len == sl.stop - sl.start
        assert splits is None or (sum(splits) == n and max(splits) - min(splits) <= 1)
    assert cover == list(range(n))


@settings(max_examples=60, deadline=None)
@given(lens=_bags, world=st.integers(min_value=1, max_value=8), seed=st.integers(0, 2**31 - 1))
def test_rowwise_filter_partitions_every_bag(lens, world, seed):
    """cloudtik_b200.rowwise.filter_owned over all ranks: every index lands on exactly one rank, in bag order, and
    the per-rank bag lengths add up to the original ones."""
    from cloudtik_b200.rowwise import filter_owned
    rs = np.random.RandomState(seed)
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    idx = rs.randint(0, 1000, size=int(sum(lens))).astype(np.int64)
    total = np.zeros(len(lens), dtype=np.int64)
    seen = []
    for r in range(world):
        li, lo = filter_owned(torch.tensor(idx), torch.tensor(off), world, r)
        total += np.diff(lo.numpy())
        seen += (li.numpy() * world + r).tolist()
    np.testing.assert_array_equal(total, np.array(lens))
    assert sorted(seen) == sorted(idx.tolist())


@settings(max_examples=40, deadline=None)
@given(n=st.integers(min_value=1, max_value=64), mir=st.sampled_from([0, 1, 7, 1000, 40_000_000]),
       seed=st.integers(0, 2**31 - 1))
def test_criteo_transform_oracle_invariants(n, mir, seed):
    rs = np.random.RandomState(seed)
    rec = rs.randint(-2**31, 2**31 - 1, size=(n, 40), dtype=np.int64).astype(np.int32)
    X, lS_o, lS_i, Y = ref.criteo_transform(rec, mir)
    assert X.shape == (n, 13) and lS_i.shape == (26, n) and lS_o.shape == (26, n) and Y.shape == (n, 1)
    np.testing.assert_array_equal(lS_o, np.tile(np.arange(n), (26, 1)))
    if mir > 0:
        assert lS_i.min() >= 0 and lS_i.max() < mir
        np.testing.assert_array_equal(lS_i, (rec[:, 14:].astype(np.int64) % mir).T)   # Python-style floor mod
    else:
        np.testing.assert_array_equal(lS_i, rec[:, 14:].astype(np.int64).T)
