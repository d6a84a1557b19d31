"""Full-size runs of the hot path (BASELINE config 2: 26 tables x 1M rows, D=128; config-3 batch 32768)
checked through size-independent properties instead of the element-by-element CPU oracle:

* integer-valued data makes every fp32 sum exact, so the pooled forward and the fused sparse update must
  equal an independent evaluation (torch index ops on the same device) BIT FOR BIT;
* the radix sort / dedup of 6.8 M keys: sortedness, stability, multiset checksums, run bookkeeping;
* interaction at B = 32768: against a bmm evaluation of the same bf16 inputs within the bf16 bound, and
  the pass-through / padding columns exactly.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

T, ROWS, D = 26, 1_000_000, 128


def _ops():
    from cloudtik_b200 import ops
    return ops


@pytest.fixture(scope="module")
def tables():
    """Integer-valued fp32 tables: W_t[r, d] = ((r * 7 + t) % 61) - 30 + (d % 5)  (|values| <= 34)."""
    r = torch.arange(ROWS, device="cuda", dtype=torch.int64)[:, None]
    d = torch.arange(D, device="cuda", dtype=torch.int64)[None, :]
    return [(((r * 7 + t) % 61) - 30 + (d % 5)).float().contiguous() for t in range(T)]


@pytest.mark.parametrize("B,P", [(262144, 1), (8192, 32), (2048, 128)])
def test_forward_full_size_exact(tables, B, P):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(B + P)
    idx = torch.randint(0, ROWS, (T, B, P), device="cuda", generator=g)
    off = torch.arange(T * B + 1, device="cuda") * P
    out = torch.empty(B, T + 1, D, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.emb_fwd(ops.TableSet(tables), idx.reshape(-1), off, B, out[:, 1:], D, (T + 1) * D, err)
    assert int(err.item()) == 0
    for t in (0, 7, 25):  # independent evaluation of three tables (all 26 would only repeat the same code path)
        want = tables[t][idx[t].reshape(-1)].view(B, P, D).sum(1)
        assert torch.equal(out[:, 1 + t], want)
    # checksum of checksums over every table: sum of all pooled values == sum over looked-up rows
    tot = sum(float(tables[t].double().sum(1)[idx[t].reshape(-1)].sum()) for t in range(T))
    assert float(out[:, 1:].double().sum()) == tot


def test_sort_dedup_full_size_properties():
    ops = _ops()
    n, bound = 26 * 262144, 26 * ROWS
    g = torch.Generator(device="cuda").manual_seed(5)
    keys = torch.randint(0, bound, (n,), device="cuda", generator=g, dtype=torch.int64)
    k32 = keys.to(torch.int32)  # < 2^31 here
    vals = torch.arange(n, device="cuda", dtype=torch.int32)
    sk, sv, uq, seg, U = ops.sort_dedup(k32, vals, bound)
    sk64, sv64 = sk.long(), sv.long()
    assert bool((sk64[1:] >= sk64[:-1]).all())                                   # sorted
    assert torch.equal(keys[sv64], sk64)                                         # sv is the permutation applied
    assert torch.equal(torch.sort(sv64).values, torch.arange(n, device="cuda"))  # ... of all positions
    same = sk64[1:] == sk64[:-1]
    assert bool((sv64[1:][same] > sv64[:-1][same]).all())                        # stable
    assert U == int(torch.unique(keys).numel())
    assert bool((uq.long()[1:] > uq.long()[:-1]).all())                          # distinct rows, ascending
    assert int(seg[0]) == 0 and int(seg[-1]) == n and bool((seg.long()[1:] > seg.long()[:-1]).all())
    assert torch.equal(sk64[seg.long()[:-1]], uq.long())                         # every run starts at its row id


@pytest.mark.parametrize("B,P", [(262144, 1), (8192, 32)])
def test_update_full_size_exact(tables, B, P):
    """lr = 1 and small-integer gradients: every partial sum is an exact fp32 integer, so the fused
    sort/dedup/segment-reduce/update must equal index_add_ exactly, duplicates and all."""
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(B * 3 + P)
    idx = torch.randint(0, ROWS, (T, B, P), device="cuda", generator=g)
    idx[3] %= 7      # a 7-row hot spot: runs of ~B*P/7 duplicates
    idx[11] %= 5000  # moderate duplicates
    off = torch.arange(T * B + 1, device="cuda") * P
    grad = torch.randint(-3, 4, (B, T + 1, D), device="cuda", generator=g).float()
    W = [w.clone() for w in tables]
    ts = ops.TableSet(W)
    ws = torch.empty(ops.emb_bwd_workspace_bytes(T * B * P, D), dtype=torch.uint8, device="cuda")
    ops.emb_bwd_update(ts, idx.reshape(-1), off, B, grad[:, 1:], D, (T + 1) * D, lr=1.0, ws=ws)
    for t in (0, 3, 11, 25):
        want = tables[t].clone()
        want.index_add_(0, idx[t].reshape(-1), -grad[:, 1 + t].repeat_interleave(P, dim=0))
        assert torch.equal(W[t], want), f"table {t}"
    # idempotence of the pair (update with g, then with -g) on integer data: back to the start, all tables
    ops.emb_bwd_update(ts, idx.reshape(-1), off, B, (-grad)[:, 1:], D, (T + 1) * D, lr=1.0, ws=ws)
    for t in range(T):
        assert torch.equal(W[t], tables[t])


def test_interaction_full_batch_vs_bmm():
    ops = _ops()
    B, F = 32768, 27
    g = torch.Generator(device="cuda").manual_seed(9)
    Tt = torch.randn(B, F, D, device="cuda", generator=g).bfloat16()
    R = ops.interact_fwd(Tt, pad_to=8)
    W = D + F * (F - 1) // 2
    assert R.shape == (B, 480) and torch.equal(R[:, :D], Tt[:, 0]) and bool((R[:, W:] == 0).all())
    Z = torch.bmm(Tt.float(), Tt.float().transpose(1, 2))
    li, lj = torch.tril_indices(F, F, offset=-1, device="cuda")
    want = Z[:, li, lj]
    assert torch.allclose(R[:, D:W].float(), want, rtol=1e-2, atol=1e-2 * D ** 0.5)
    dR = torch.zeros(B, 480, dtype=torch.bfloat16, device="cuda")
    dR[:, :W] = torch.randn(B, W, device="cuda", generator=g).bfloat16()
    dT = ops.interact_bwd(Tt, dR)
    dZ = torch.zeros(B, F, F, device="cuda")
    dZ[:, li, lj] = dR[:, D:W].float()
    wdT = torch.bmm(dZ + dZ.transpose(1, 2), Tt.float())
    wdT[:, 0] += dR[:, :D].float()
    assert torch.allclose(dT.float(), wdT, rtol=1e-2, atol=1e-2 * F ** 0.5)
