"""Device input pipeline: b200_criteo_transform / cloudtik_b200.criteo_input against
the reference-generated fixture and the numpy oracle -- ids, offsets and labels bit-exact, log(x+1) <= 1e-6 rel."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _check(got, want_X, want_lS_o, want_lS_i, want_Y):
    X, lS_o, lS_i, Y = got
    assert torch.equal(lS_i.cpu(), torch.from_numpy(want_lS_i))
    assert torch.equal(lS_o.cpu(), torch.from_numpy(want_lS_o))
    assert torch.equal(Y.cpu(), torch.from_numpy(want_Y))
    np.testing.assert_allclose(X.cpu().numpy(), want_X, rtol=1e-6, atol=0)


def test_transform_features_matches_reference_fixture(golden):
    from cloudtik_b200.criteo_input import _transform_features
    g = golden("criteo_transform")
    for tag in "abc":
        rec = torch.from_numpy(g[f"{tag}_records"]).cuda()
        got = _transform_features(x_int_batch=rec[:, 1:14], x_cat_batch=rec[:, 14:], y_batch=rec[:, 0],
                                  max_ind_range=int(g[f"{tag}_max_ind_range"]), flag_input_torch_tensor=True)
        _check(got, g[f"{tag}_X"], g[f"{tag}_lS_o"], g[f"{tag}_lS_i"], g[f"{tag}_Y"])


@pytest.mark.parametrize("N", [1, 63, 64, 65, 4099, 32768])
def test_transform_vs_oracle_sizes_slices_and_padding(N):
    from cloudtik_b200 import ops
    from oracle import ops_numpy as ref
    rs = np.random.RandomState(N)
    rec = rs.randint(-2**31, 2**31 - 1, size=(N, 40), dtype=np.int64).astype(np.int32)
    rec[:, 0] = rs.randint(0, 2, size=N)
    rec[:, 1:14] = rs.randint(0, 1 << 20, size=(N, 13))
    mir = 40_000_000
    wX, wo, wi, wY = ref.criteo_transform(rec, mir)
    d = torch.from_numpy(rec).cuda()
    X, Y, ind, off = ops.criteo_transform(d, mir)
    assert torch.equal(ind.cpu().view(26, N), torch.from_numpy(wi))
    assert torch.equal(off.cpu(), torch.arange(26 * N + 1))
    assert torch.equal(Y.cpu(), torch.from_numpy(wY))
    np.testing.assert_allclose(X.cpu().numpy(), wX, rtol=1e-6)
    # a rank's view: tables [7, 12) for every sample, dense rows of the second half, K padded to 16 columns
    r0 = N // 2
    X2, Y2, ind2, off2 = ops.criteo_transform(d, mir, tab_begin=7, tab_count=5, row_begin=r0, row_count=N - r0,
                                              dense_stride=16)
    assert torch.equal(ind2.cpu().view(5, N), torch.from_numpy(wi[7:12]))
    assert torch.equal(off2.cpu(), torch.arange(5 * N + 1))
    assert torch.equal(Y2.cpu(), torch.from_numpy(wY[r0:]))
    np.testing.assert_allclose(X2[:, :13].cpu().numpy(), wX[r0:], rtol=1e-6)
    assert bool((X2[:, 13:] == 0).all())


def test_bin_dataset_round_trip(tmp_path):
    """numpy_to_binary -> CriteoBinDataset (same file format as the reference) -> model inputs, incl. the short last batch."""
    from cloudtik_b200.criteo_input import CriteoBinDataset, numpy_to_binary
    from oracle import ops_numpy as ref
    rs = np.random.RandomState(3)
    N, bs = 1000, 256
    y = rs.randint(0, 2, size=N)
    x_int = rs.randint(0, 5000, size=(N, 13))
    x_cat = rs.randint(0, 2**31 - 1, size=(N, 26))
    f = tmp_path / "day_0.bin"
    rec = numpy_to_binary(y, x_int, x_cat, str(f))
    ds = CriteoBinDataset(str(f), None, batch_size=bs, max_ind_range=10_000_000)
    assert len(ds) == 4
    seen = 0
    for i in range(len(ds)):
        X, lS_o, lS_i, Y = ds[i]
        want = ref.criteo_transform(rec[i * bs:(i + 1) * bs], 10_000_000)
        _check((X, lS_o, lS_i, Y), want[0], want[1], want[2], want[3])
        seen += X.shape[0]
    assert seen == N
    # merged form feeds MergedEmbeddingBagWithSGD directly: (X, offsets, indices, Y)
    dm = CriteoBinDataset(str(f), None, batch_size=bs, max_ind_range=10_000_000, merged=True)
    X, off, ind, Y = dm[1]
    want = ref.criteo_transform(rec[bs:2 * bs], 10_000_000)
    assert torch.equal(ind.cpu().view(26, bs), torch.from_numpy(want[2])) and off.numel() == 26 * bs + 1


def test_log_of_negative_counts_follows_torch():
    """x = -1 -> -inf, x < -1 -> nan: same as torch.log(x.float() + 1); nothing is clamped."""
    from cloudtik_b200 import ops
    rec = torch.zeros(4, 40, dtype=torch.int32)
    rec[:, 1] = torch.tensor([-1, -5, 0, 7])
    X = ops.criteo_transform(rec.cuda(), 0)[0].cpu()
    want = torch.log(rec[:, 1].float() + 1)
    assert X[0, 0] == want[0] and torch.isnan(X[1, 0]) and X[2, 0] == 0 and abs(float(X[3, 0] - want[3])) < 1e-6
