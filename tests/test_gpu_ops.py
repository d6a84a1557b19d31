"""GPU parity tests of the C-ABI kernels against the CPU oracle (oracle/ops_numpy.py).

Tolerances (BASELINE.json north_star): bit-exact for index sort / dedup / bucketing;
<= 1e-5 relative for fp32 sums / updates, <= 1e-2 for bf16.
"""
import numpy as np
import pytest
import torch

from oracle import ops_numpy as ref

pytestmark = pytest.mark.gpu

RTOL32, ATOL32 = 1e-5, 1e-6
RTOL16, ATOL16 = 1e-2, 1e-2


def _ops():
    from cloudtik_b200 import ops
    return ops


def _make_bags(rs, rows, B, pooling, ragged):
    """Per-table (indices, offsets) lists; ragged => random lengths in [0, 2*pooling]."""
    idx_l, off_l = [], []
    for n in rows:
        if ragged:
            lens = rs.randint(0, 2 * pooling + 1, size=B)
        else:
            lens = np.full(B, pooling)
        off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
        idx_l.append(rs.randint(0, n, size=int(lens.sum())).astype(np.int64))
        off_l.append(off)
    return idx_l, off_l


def _tables(rs, rows, D, dtype):
    Ws = [rs.uniform(-1, 1, size=(n, D)).astype(np.float32) for n in rows]
    tw = [torch.tensor(W).to(dtype).cuda() for W in Ws]
    Wr = [t.float().cpu().numpy() for t in tw]  # the values actually stored (bf16-rounded)
    return Wr, tw


FWD_CASES = [
    # rows, B, D, pooling, ragged, wdtype, odtype
    ([1000, 37, 5000], 257, 128, 1, False, torch.float32, torch.float32),
    ([1000, 37, 5000], 300, 128, 6, True, torch.float32, torch.float32),
    ([64, 64], 130, 16, 1, False, torch.float32, torch.float32),
    ([64, 64], 33, 8, 3, True, torch.float32, torch.float32),
    ([50, 20, 10], 7, 2, 2, True, torch.float32, torch.float32),      # scalar path
    ([500, 700], 129, 64, 1, False, torch.bfloat16, torch.bfloat16),
    ([500, 700], 64, 128, 40, True, torch.bfloat16, torch.float32),
    ([500], 100, 128, 4, True, torch.float32, torch.bfloat16),
    ([300, 200], 50, 256, 2, True, torch.float32, torch.float32),      # NV=2
    ([300], 40, 512, 1, False, torch.float32, torch.float32),          # NV=4
    ([2000], 64, 128, 200, False, torch.float32, torch.float32),       # long bags, unstaged indices
    ([100, 100], 5, 24, 2, True, torch.float32, torch.float32),        # V=6 (non power of two)
]


@pytest.mark.parametrize("rows,B,D,pooling,ragged,wdt,odt", FWD_CASES)
def test_emb_fwd(rows, B, D, pooling, ragged, wdt, odt):
    ops = _ops()
    rs = np.random.RandomState(len(rows) * 1000 + B + D)
    Wr, tw = _tables(rs, rows, D, wdt)
    idx_l, off_l = _make_bags(rs, rows, B, pooling, ragged)
    indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    ts = ops.TableSet(tw)
    T = len(rows)
    F = T + 1
    # strided [B, F, D] layout (slot 0 reserved for the dense feature) ...
    buf = torch.full((B, F, D), 7.0, dtype=odt, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.emb_fwd(ts, torch.tensor(indices).cuda(), torch.tensor(offsets).cuda(), B, buf[:, 1:], D, F * D, err)
    # ... and table-major contiguous [T, B, D]
    out2 = torch.empty((T, B, D), dtype=odt, device="cuda")
    ops.emb_fwd(ts, torch.tensor(indices).cuda(), torch.tensor(offsets).cuda(), B, out2, B * D, D, err)
    torch.cuda.synchronize()
    assert int(err.item()) == 0
    got = buf.float().cpu().numpy()
    got2 = out2.float().cpu().numpy()
    assert (got[:, 0] == 7.0).all(), "slot 0 must be untouched"
    rtol, atol = (RTOL32, ATOL32) if odt == torch.float32 else (RTOL16, ATOL16)
    for t in range(T):
        want = ref.embedding_bag_sum(Wr[t], idx_l[t], off_l[t])
        np.testing.assert_allclose(got[:, 1 + t], want, rtol=rtol, atol=atol * pooling)
        np.testing.assert_array_equal(got[:, 1 + t], got2[t])


def test_emb_fwd_flags_bad_index():
    ops = _ops()
    tw = [torch.ones(10, 128, device="cuda")]
    ts = ops.TableSet(tw)
    idx = torch.tensor([1, 99, 2], device="cuda")
    off = torch.tensor([0, 1, 2, 3], device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    out = torch.empty(3, 128, device="cuda")
    ops.emb_fwd(ts, idx, off, 3, out, 3 * 128, 128, err)
    torch.cuda.synchronize()
    assert int(err.item()) == 1
    assert out[0].eq(1).all() and out[1].eq(0).all() and out[2].eq(1).all()


def _i32(a):
    return torch.tensor(np.asarray(a, dtype=np.uint32).view(np.int32)).cuda()


def _u32(t):
    return t.cpu().numpy().view(np.uint32).astype(np.int64)


@pytest.mark.parametrize("n,bound", [(0, 10), (1, 1), (31, 7), (4096, 200), (4097, 70000), (100003, 1 << 20),
                                     (50000, 187_000_000), (70000, 1 << 32), (300000, 1000)])
def test_sort_dedup_bit_exact(n, bound):
    ops = _ops()
    rs = np.random.RandomState(n % 9973 + 1)
    keys = rs.randint(0, bound, size=n, dtype=np.int64)
    vals = np.arange(n, dtype=np.int64)
    sk, sv, uq, seg, U = ops.sort_dedup(_i32(keys), _i32(vals), bound)
    wk, wperm, wuq, wcnt, wseg = ref.sort_dedup(keys)
    np.testing.assert_array_equal(_u32(sk), wk)
    np.testing.assert_array_equal(_u32(sv), wperm)   # payload = position => the stable permutation
    assert U == len(wuq)
    np.testing.assert_array_equal(_u32(uq), wuq)
    np.testing.assert_array_equal(_u32(seg), wseg)
    np.testing.assert_array_equal(np.diff(_u32(seg)), wcnt)


def test_sort_dedup_golden_kat(golden):
    ops = _ops()
    g = golden("index_kats")
    keys = g["idx"]
    sk, sv, uq, seg, U = ops.sort_dedup(_i32(keys), _i32(np.arange(len(keys))), 10)
    np.testing.assert_array_equal(_u32(sk), g["sorted"])
    np.testing.assert_array_equal(_u32(sv), g["perm"])
    np.testing.assert_array_equal(_u32(uq), g["coalesced_indices"])
    np.testing.assert_array_equal(np.diff(_u32(seg)), g["counts"])


BWD_CASES = [
    # rows, B, D, pooling, ragged, wdtype, gdtype, opt
    ([1000, 37, 5000], 257, 128, 1, False, torch.float32, torch.float32, "sgd"),
    ([50, 37, 20], 300, 128, 6, True, torch.float32, torch.float32, "sgd"),       # heavy duplicates
    ([64, 64], 130, 16, 1, False, torch.float32, torch.float32, "sgd"),
    ([50, 20, 10], 7, 2, 2, True, torch.float32, torch.float32, "sgd"),           # scalar path
    ([500, 700], 129, 64, 2, True, torch.bfloat16, torch.bfloat16, "sgd"),
    ([500, 700], 129, 128, 2, True, torch.float32, torch.bfloat16, "sgd"),
    ([500], 100, 128, 3, True, torch.bfloat16, torch.float32, "sgd"),
    ([300, 200], 50, 256, 2, True, torch.float32, torch.float32, "sgd"),
    ([1000, 37, 5000], 257, 128, 4, True, torch.float32, torch.float32, "adagrad"),
    ([40, 30], 64, 8, 3, True, torch.float32, torch.float32, "adagrad"),
    ([50, 20, 10], 7, 2, 2, True, torch.float32, torch.float32, "adagrad"),
    ([40, 300], 70, 1024, 3, True, torch.bfloat16, torch.bfloat16, "sgd"),        # 2 KB rows: 64 KB of in-block pieces
    ([40, 300], 70, 512, 3, True, torch.float32, torch.float32, "sgd"),           # widest fp32 row (V = 128)
]


@pytest.mark.parametrize("rows,B,D,pooling,ragged,wdt,gdt,opt", BWD_CASES)
def test_emb_bwd_update(rows, B, D, pooling, ragged, wdt, gdt, opt):
    ops = _ops()
    rs = np.random.RandomState(len(rows) * 77 + B + D)
    Wr, tw = _tables(rs, rows, D, wdt)
    idx_l, off_l = _make_bags(rs, rows, B, pooling, ragged)
    indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    T = len(rows)
    F = T + 1
    g = torch.tensor(rs.randn(B, F, D).astype(np.float32)).to(gdt).cuda()
    gr = g.float().cpu().numpy()
    ts = ops.TableSet(tw)
    lr = 0.3
    states_np = None
    if opt == "adagrad":
        st = [torch.tensor(rs.rand(n, D).astype(np.float32)).cuda() for n in rows]
        states_np = [s.cpu().numpy().copy() for s in st]
        ts.set_states(st)
    ws = torch.empty(ops.emb_bwd_workspace_bytes(len(indices), D), dtype=torch.uint8, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    lr_t = torch.tensor([lr], device="cuda")
    ops.emb_bwd_update(ts, torch.tensor(indices).cuda(), torch.tensor(offsets).cuda(), B, g[:, 1:], D, F * D,
                       lr=123.0, ws=ws, optimizer=ops.OPT_ADAGRAD if opt == "adagrad" else ops.OPT_SGD,
                       eps=1e-10, lr_tensor=lr_t, error=err)  # device lr must win over the by-value one
    torch.cuda.synchronize()
    assert int(err.item()) == 0
    rtol, atol = (RTOL32, ATOL32 * 10) if wdt == torch.float32 else (RTOL16, ATOL16)
    for t in range(T):
        W = Wr[t].copy()
        if opt == "adagrad":
            ref.sparse_adagrad_update(W, states_np[t], idx_l[t], off_l[t], gr[:, 1 + t], lr, 1e-10)
            np.testing.assert_allclose(ts.states[t].cpu().numpy(), states_np[t], rtol=1e-5, atol=1e-6)
        else:
            ref.sparse_sgd_update(W, idx_l[t], off_l[t], gr[:, 1 + t], lr)
        np.testing.assert_allclose(tw[t].float().cpu().numpy(), W, rtol=rtol, atol=atol)


HOT_CASES = [
    # rows, B, D, pooling, wdtype, gdtype, opt  -- runs far longer than the 128-position hot window
    ([3, 500, 7], 3000, 128, 1, torch.float32, torch.float32, "sgd"),
    ([3, 40], 1500, 64, 2, torch.bfloat16, torch.bfloat16, "sgd"),
    ([4, 100000], 700, 128, 1, torch.float32, torch.bfloat16, "sgd"),
    ([4], 700, 128, 1, torch.float32, torch.float32, "adagrad"),
    ([2, 9], 1111, 16, 3, torch.float32, torch.float32, "sgd"),
]


@pytest.mark.parametrize("rows,B,D,pooling,wdt,gdt,opt", HOT_CASES)
def test_emb_bwd_hot_rows(rows, B, D, pooling, wdt, gdt, opt):
    """Hot rows (thousands of duplicates: the tiny Criteo tables) go through the windowed
    partial-sum path; checked against a float64 evaluation of the same update."""
    ops = _ops()
    rs = np.random.RandomState(B + D)
    Wr, tw = _tables(rs, rows, D, wdt)
    idx_l, off_l = _make_bags(rs, rows, B, pooling, False)
    indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    T, F = len(rows), len(rows) + 1
    g = torch.tensor(rs.randn(B, F, D).astype(np.float32)).to(gdt).cuda()
    gr = g.float().cpu().numpy().astype(np.float64)
    ts = ops.TableSet(tw)
    lr = 0.01
    if opt == "adagrad":
        ts.set_states([torch.zeros(n, D, device="cuda") for n in rows])
    ws = torch.empty(ops.emb_bwd_workspace_bytes(len(indices), D), dtype=torch.uint8, device="cuda")
    ops.emb_bwd_update(ts, torch.tensor(indices).cuda(), torch.tensor(offsets).cuda(), B, g[:, 1:], D, F * D,
                       lr=lr, ws=ws, optimizer=ops.OPT_ADAGRAD if opt == "adagrad" else ops.OPT_SGD)
    torch.cuda.synchronize()
    for t in range(T):
        G = np.zeros((rows[t], D))
        np.add.at(G, idx_l[t], gr[:, 1 + t][ref.offset2bag(off_l[t], len(idx_l[t]))])
        if opt == "adagrad":
            want = Wr[t] - lr * G / (np.sqrt(G * G) + 1e-10)
        else:
            want = Wr[t] - lr * G
        got = tw[t].float().cpu().numpy()
        if wdt == torch.float32:
            np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5 * np.abs(lr * G).max())
        else:
            np.testing.assert_allclose(got, want, rtol=1e-2, atol=1e-2)


def test_emb_bwd_hot_window_boundaries():
    """Crafted run lengths around the window size (127, 128, 129, 256, 257, 1, 640...) in one table."""
    ops = _ops()
    D, lens = 128, [127, 128, 129, 1, 256, 257, 1, 1, 640, 5, 383, 130]
    rows = [len(lens) + 3]
    idx = np.concatenate([np.full(n, r) for r, n in enumerate(lens)]).astype(np.int64)
    rs = np.random.RandomState(0)
    rs.shuffle(idx)
    B = len(idx)
    off = np.arange(B + 1, dtype=np.int64)
    W = torch.zeros(rows[0], D, device="cuda")
    g = torch.tensor(rs.randint(-3, 4, size=(B, D)).astype(np.float32)).cuda()  # integers: sums are exact
    ts = ops.TableSet([W])
    ws = torch.empty(ops.emb_bwd_workspace_bytes(B, D), dtype=torch.uint8, device="cuda")
    ops.emb_bwd_update(ts, torch.tensor(idx).cuda(), torch.tensor(off).cuda(), B, g, B * D, D, lr=1.0, ws=ws)
    torch.cuda.synchronize()
    want = np.zeros((rows[0], D), np.float32)
    np.add.at(want, idx, -g.cpu().numpy())
    np.testing.assert_array_equal(W.cpu().numpy(), want)


def test_emb_bwd_reader_right_behind_update_with_all_distinct_rows():
    """No run crosses any block border here (every key is distinct), so no group of the join grid has rows to
    add -- the join grid must still order itself behind the window kernel (programmatic dependent launch), or the
    next kernel on the stream (here: a copy of the table) could read rows that are still being written."""
    ops = _ops()
    D, n = 128, 1 << 20
    rs = np.random.RandomState(5)
    idx = rs.permutation(n).astype(np.int64)
    off = np.arange(n + 1, dtype=np.int64)
    ti, to = torch.tensor(idx).cuda(), torch.tensor(off).cuda()
    g = torch.ones(n, D, device="cuda")
    ws = torch.empty(ops.emb_bwd_workspace_bytes(n, D), dtype=torch.uint8, device="cuda")
    for rep in range(3):
        W = torch.zeros(n, D, device="cuda")
        ts = ops.TableSet([W])
        ops.emb_bwd_prepare(ts, ti, to, n, ws)
        torch.cuda.synchronize()
        ops.emb_bwd_apply(ts, n, n, g, n * D, D, 1.0, ws)
        snap = W.clone()  # launched right behind the update on the same stream
        torch.cuda.synchronize()
        assert bool((snap == -1.0).all()), f"rep {rep}: {int((snap != -1.0).sum())} stale elements"


@pytest.mark.parametrize("rows,B,D,pooling,ragged,gdt", [([3, 1543, 63, 10], 3000, 128, 1, False, torch.float32),
                                                         ([3, 40, 7], 700, 128, 3, True, torch.bfloat16),
                                                         ([5, 9], 64, 16, 2, True, torch.float32),
                                                         ([50, 20, 10], 7, 2, 2, True, torch.float32)])
def test_emb_bwd_dense_grad(rows, B, D, pooling, ragged, gdt):
    """Dense gradient of small (replicated) tables: prepare + b200_emb_bwd_dense_grad vs the oracle's scatter-add
    (what autograd gives nn.EmbeddingBag(sparse=False), T/dlrm_s_pytorch.py:266-269); untouched rows stay zero."""
    ops = _ops()
    rs = np.random.RandomState(B + D)
    _, tw = _tables(rs, rows, D, torch.float32)
    idx_l, off_l = _make_bags(rs, rows, B, pooling, ragged)
    indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    T, F = len(rows), len(rows) + 1
    g = torch.tensor(rs.randn(B, F, D).astype(np.float32)).to(gdt).cuda()
    gr = g.float().cpu().numpy().astype(np.float64)
    ts = ops.TableSet(tw)
    ws = torch.empty(ops.emb_bwd_workspace_bytes(len(indices), D), dtype=torch.uint8, device="cuda")
    ops.emb_bwd_prepare(ts, torch.tensor(indices).cuda(), torch.tensor(offsets).cuda(), B, ws)
    G = torch.zeros(sum(rows), D, device="cuda")
    ops.emb_bwd_dense_grad(ts, len(indices), B, g[:, 1:], D, F * D, G, ws)
    got = G.cpu().numpy()
    base = 0
    for t in range(T):
        want = np.zeros((rows[t], D))
        np.add.at(want, idx_l[t], gr[:, 1 + t][ref.offset2bag(off_l[t], len(idx_l[t]))])
        np.testing.assert_allclose(got[base:base + rows[t]], want, rtol=1e-5, atol=1e-5 * max(1.0, np.abs(want).max()))
        base += rows[t]


@pytest.mark.parametrize("opt", ["sgd", "adagrad"])
def test_bf16_tables_with_trail_update_in_fp32(opt):
    """bf16 tables + 16-bit trail (the reference's split-SGD storage): the fused update must behave like
    an fp32 update -- tiny steps that a plain bf16 row would round away accumulate exactly."""
    ops = _ops()
    rs = np.random.RandomState(12)
    rows, B, D, lr = [3, 900, 50], 700, 128, 1e-4
    Ws = [rs.uniform(-1, 1, size=(n, D)).astype(np.float32) for n in rows]
    his, los = [], []
    for W in Ws:
        bits = torch.tensor(W).view(torch.int32)
        his.append((bits >> 16).to(torch.int16).view(torch.bfloat16).cuda().contiguous())
        los.append((bits & 0xFFFF).to(torch.int16).cuda().contiguous())
    ts = ops.TableSet(his)
    ts.set_trails(los)
    states = None
    if opt == "adagrad":
        ts.set_states([torch.zeros(n, D, device="cuda") for n in rows])
        states = [np.zeros((n, D), np.float32) for n in rows]
    T, F = len(rows), len(rows) + 1
    Wref = [W.copy() for W in Ws]
    ws = None
    for step in range(3):
        idx_l, off_l = _make_bags(rs, rows, B, 2, True)
        indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
        g = torch.tensor(rs.randn(B, F, D).astype(np.float32)).cuda()
        ws = torch.empty(ops.emb_bwd_workspace_bytes(len(indices), D), dtype=torch.uint8, device="cuda")
        ops.emb_bwd_update(ts, torch.tensor(indices).cuda(), torch.tensor(offsets).cuda(), B, g[:, 1:], D, F * D,
                           lr=lr, ws=ws, optimizer=ops.OPT_ADAGRAD if opt == "adagrad" else ops.OPT_SGD)
        gr = g.cpu().numpy()
        for t in range(T):
            if opt == "adagrad":
                ref.sparse_adagrad_update(Wref[t], states[t], idx_l[t], off_l[t], gr[:, 1 + t], lr, 1e-10)
            else:
                G = np.zeros((rows[t], D))
                np.add.at(G, idx_l[t], gr[:, 1 + t][ref.offset2bag(off_l[t], len(idx_l[t]))].astype(np.float64))
                Wref[t] = (Wref[t] - lr * G).astype(np.float32)
    torch.cuda.synchronize()
    for t in range(T):
        hi = his[t].view(torch.int16).to(torch.int32) << 16
        lo = los[t].to(torch.int32) & 0xFFFF
        got = (hi | lo).view(torch.float32).cpu().numpy()
        np.testing.assert_allclose(got, Wref[t], rtol=2e-5, atol=1e-6)
        # a plain bf16 row could not have followed: the accumulated change is below bf16 resolution
        moved = np.abs(got - Ws[t]) > 0
        assert moved.any() and np.abs(got - Ws[t])[moved].min() < 1e-3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_interaction_golden(golden, dtype):
    ops = _ops()
    g = golden("interaction")
    for tag in "abc":
        x, ly = g[f"{tag}_x"], g[f"{tag}_ly"]
        T = torch.tensor(np.concatenate([x[:, None], ly.transpose(1, 0, 2)], axis=1)).to(dtype).cuda().contiguous()
        R = ops.interact_fwd(T)
        dR = torch.tensor(g[f"{tag}_dR"]).to(dtype).cuda()
        dT = ops.interact_bwd(T, dR)
        torch.cuda.synchronize()
        if dtype == torch.float32:
            np.testing.assert_allclose(R.cpu().numpy(), g[f"{tag}_R"], rtol=1e-5, atol=1e-5)
            np.testing.assert_allclose(dT[:, 0].cpu().numpy(), g[f"{tag}_dx"], rtol=1e-5, atol=5e-5)
            np.testing.assert_allclose(dT[:, 1:].cpu().numpy().transpose(1, 0, 2), g[f"{tag}_dly"], rtol=1e-5, atol=5e-5)
        else:
            Tf = T.float().cpu().numpy()
            want = ref.interact_fwd(Tf[:, 0], [Tf[:, k] for k in range(1, Tf.shape[1])])
            scale = np.abs(want).max()
            np.testing.assert_allclose(R.float().cpu().numpy(), want, rtol=1e-2, atol=1e-2 * scale)
            dx, dly = ref.interact_bwd(Tf[:, 0], [Tf[:, k] for k in range(1, Tf.shape[1])], dR.float().cpu().numpy())
            wdT = np.stack([dx] + dly, axis=1)
            np.testing.assert_allclose(dT.float().cpu().numpy(), wdT, rtol=1e-2, atol=1e-2 * np.abs(wdT).max())


@pytest.mark.parametrize("B,F,D,pad", [(1, 27, 128, 1), (3, 27, 128, 8), (4, 27, 128, 8), (5, 2, 128, 1),
                                       (1000, 27, 128, 8), (613, 32, 128, 1), (77, 27, 256, 8), (2048, 9, 256, 1),
                                       (4099, 27, 128, 8)])
def test_interaction_tensor_core_path(B, F, D, pad):
    """bf16, F <= 32, D in {128, 256}: the tcgen05 kernels.  Checked against a float64 evaluation of
    the same bf16 inputs; outputs are bf16-rounded, so <= 1e-2 relative (north_star bf16 bound)."""
    ops = _ops()
    rs = np.random.RandomState(B + F + D)
    T = torch.tensor(rs.randn(B, F, D).astype(np.float32)).bfloat16().cuda()
    Tf = T.float().cpu().numpy().astype(np.float64)
    W = D + F * (F - 1) // 2
    R = ops.interact_fwd(T, pad_to=pad)
    ldr = R.shape[1]
    assert ldr % pad == 0 and ldr >= W
    want = ref.interact_fwd(Tf[:, 0], [Tf[:, k] for k in range(1, F)])
    got = R.float().cpu().numpy()
    np.testing.assert_allclose(got[:, :W], want, rtol=1e-2, atol=1e-2 * np.sqrt(D))
    np.testing.assert_array_equal(got[:, :D], Tf[:, 0])          # pass-through is exact
    assert (got[:, W:] == 0).all()
    dR = torch.zeros(B, ldr, dtype=torch.bfloat16, device="cuda")
    dR[:, :W] = torch.tensor(rs.randn(B, W).astype(np.float32)).bfloat16().cuda()
    dT = ops.interact_bwd(T, dR)
    torch.cuda.synchronize()
    dRf = dR.float().cpu().numpy().astype(np.float64)[:, :W]
    dx, dly = ref.interact_bwd(Tf[:, 0], [Tf[:, k] for k in range(1, F)], dRf)
    wdT = np.stack([dx] + dly, axis=1)
    np.testing.assert_allclose(dT.float().cpu().numpy(), wdT, rtol=1e-2, atol=1e-2 * np.sqrt(F))


@pytest.mark.parametrize("B,D,pad", [(1, 128, 1), (5, 128, 8), (1000, 128, 8), (4099, 128, 1), (333, 64, 8), (777, 32, 1)])
def test_interaction_fp32_register_tiled_path(B, D, pad):
    """fp32, F = 27, D in {32, 64, 128}: the warp-per-sample register-tiled kernels (interact_f32.cu), against a
    float64 evaluation of the same inputs at the north_star fp32 bound (1e-5 relative)."""
    ops = _ops()
    F = 27
    rs = np.random.RandomState(B + D)
    Tn = rs.randn(B, F, D).astype(np.float32)
    T = torch.tensor(Tn).cuda()
    W = D + F * (F - 1) // 2
    R = ops.interact_fwd(T, pad_to=pad)
    ldr = R.shape[1]
    T64 = Tn.astype(np.float64)
    want = ref.interact_fwd(T64[:, 0], [T64[:, k] for k in range(1, F)])
    got = R.cpu().numpy()
    np.testing.assert_allclose(got[:, :W], want, rtol=1e-5, atol=1e-5 * np.sqrt(D))
    np.testing.assert_array_equal(got[:, :D], Tn[:, 0])
    assert (got[:, W:] == 0).all()
    dRn = np.zeros((B, ldr), dtype=np.float32)
    dRn[:, :W] = rs.randn(B, W)
    dT = ops.interact_bwd(T, torch.tensor(dRn).cuda())
    dx, dly = ref.interact_bwd(T64[:, 0], [T64[:, k] for k in range(1, F)], dRn[:, :W].astype(np.float64))
    wdT = np.stack([dx] + dly, axis=1)
    np.testing.assert_allclose(dT.cpu().numpy(), wdT, rtol=1e-5, atol=1e-5 * np.sqrt(F))


def test_interaction_kat():
    ops = _ops()
    T = torch.tensor([[[1.0, 2.0], [3.0, 4.0], [5.0, 6.0]]], device="cuda")
    R = ops.interact_fwd(T)
    assert R.cpu().tolist() == [[1.0, 2.0, 11.0, 17.0, 39.0]]


@pytest.mark.parametrize("B,N", [(1, 8), (127, 128), (128, 256), (1000, 1024), (4099, 512), (32768, 1024)])
def test_relu_bwd_bias_matches_torch(B, N):
    """b200_relu_bwd_bias: the masked gradient must equal aten.threshold_backward BIT FOR BIT (it only zeroes
    entries), the bias gradient must equal the fp64 column sum of that masked matrix to fp32 accuracy."""
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(B + N)
    y = torch.relu(torch.randn(B, N, device="cuda", generator=g)).bfloat16()
    y[0, 0] = float("nan")      # NaN activation keeps the gradient (torch: `self <= 0 ? 0 : grad`)
    y[-1, -1] = -0.0
    dy = torch.randn(B, N, device="cuda", generator=g).bfloat16()
    dz, db = ops.relu_bwd_bias(dy, y)
    want = torch.ops.aten.threshold_backward(dy, y, 0)
    assert torch.equal(dz.view(torch.int16), want.view(torch.int16))
    ref_db = want.double().sum(0)
    scale = want.double().abs().sum(0).clamp_min(1.0)
    assert bool(((db.double() - ref_db).abs() <= 1e-6 * scale).all())
    # deterministic: same bits on a second call
    dz2, db2 = ops.relu_bwd_bias(dy, y)
    assert torch.equal(db, db2)


@pytest.mark.parametrize("D,variant", [(128, "sgd_fp32"), (128, "sgd_bf16grad"), (128, "adagrad"), (128, "bf16_trail"),
                                       (16, "sgd_fp32"), (64, "adagrad"), (256, "sgd_fp32"), (128, "coalesce")])
def test_emb_bwd_block_and_window_borders(D, variant):
    """Crafted run lengths around the window (32 positions) and CTA (8 .. 64 windows) borders of the update kernel:
    runs that end exactly on a border, start on one, span several CTAs (in-block join, per-CTA partial rows, the
    cross-CTA join) for every update flavour.  Integer gradients: all sums are exact, so SGD must be bit-exact."""
    ops = _ops()
    lens = [31, 1, 32, 33, 255, 1, 256, 257, 3, 511, 513, 1, 1, 1024, 7, 2047, 2, 2049, 700, 5000, 9, 64, 64, 64, 1, 8191]
    n_rows = len(lens) + 2
    idx = np.concatenate([np.full(n, r) for r, n in enumerate(lens)]).astype(np.int64)
    rs = np.random.RandomState(D + len(variant))
    rs.shuffle(idx)
    B = len(idx)
    off = np.arange(B + 1, dtype=np.int64)
    gdt = torch.bfloat16 if variant in ("sgd_bf16grad", "bf16_trail") else torch.float32
    g = torch.tensor(rs.randint(-3, 4, size=(B, D)).astype(np.float32)).to(gdt).cuda()
    G = np.zeros((n_rows, D), np.float64)
    np.add.at(G, idx, g.float().cpu().numpy().astype(np.float64))
    d_idx, d_off = torch.tensor(idx).cuda(), torch.tensor(off).cuda()
    ws = torch.empty(ops.emb_bwd_workspace_bytes(B, D), dtype=torch.uint8, device="cuda")
    W0 = rs.randint(-5, 6, size=(n_rows, D)).astype(np.float32)
    if variant == "coalesce":
        ts = ops.TableSet([torch.tensor(W0).cuda()])
        rows_u, vals = ops.emb_bwd_coalesce(ts, d_idx, d_off, B, g, B * D, D, ws)
        np.testing.assert_array_equal(rows_u.cpu().numpy(), np.arange(len(lens)))
        np.testing.assert_array_equal(vals.cpu().numpy(), G[:len(lens)].astype(np.float32))
        return
    if variant == "bf16_trail":
        w = torch.tensor(W0).cuda()
        bits = w.view(torch.int32)
        trail = (bits & 0xFFFF).to(torch.int16)
        ts = ops.TableSet([(bits >> 16).to(torch.int16).view(torch.bfloat16).contiguous()])
        ts.set_trails([trail])
    else:
        ts = ops.TableSet([torch.tensor(W0).cuda()])
    if variant == "adagrad":
        ts.set_states([torch.full((n_rows, D), 0.5, device="cuda")])
    ops.emb_bwd_update(ts, d_idx, d_off, B, g, B * D, D, lr=1.0, ws=ws,
                       optimizer=ops.OPT_ADAGRAD if variant == "adagrad" else ops.OPT_SGD)
    torch.cuda.synchronize()
    if variant == "bf16_trail":
        hi = ts.weights[0].view(torch.int16).to(torch.int32) << 16
        got = (hi | (ts.trails[0].to(torch.int32) & 0xFFFF)).view(torch.float32).cpu().numpy()
    else:
        got = ts.weights[0].cpu().numpy()
    if variant == "adagrad":
        st = 0.5 + G * G
        want = W0 - G / (np.sqrt(st) + 1e-10)
        np.testing.assert_allclose(got, want, rtol=2e-6, atol=1e-6)
        np.testing.assert_allclose(ts.states[0].cpu().numpy(), st, rtol=1e-6)
    else:
        np.testing.assert_array_equal(got, (W0 - G).astype(np.float32))


@pytest.mark.parametrize("W,lN,dtype", [(1, 96, torch.float32), (2, 64, torch.float32), (4, 48, torch.bfloat16),
                                        (8, 32, torch.float32), (3, 40, torch.float32)])
def test_sharded_entry_points_on_one_gpu(W, lN, dtype):
    """b200_emb_fwd_sharded / b200_emb_bwd_apply_sharded with the W destination "ranks" emulated by W buffers on
    one GPU: the rows must land where a pooled forward followed by the reference's all-to-all would put them, and
    the update fed from the W gradient buffers must equal the plain update fed from one contiguous gradient."""
    ops = _ops()
    rs = np.random.RandomState(W * 100 + lN)
    rows, D, F, feat0 = [700, 5, 3000], 128, 9, 4   # this "rank" owns features 4, 5, 6 of 9
    T, B = len(rows), W * lN
    Wr, tw = _tables(rs, rows, D, torch.float32)
    idx_l, off_l = _make_bags(rs, rows, B, 3, True)
    indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    d_idx, d_off = torch.tensor(indices).cuda(), torch.tensor(offsets).cuda()
    ts = ops.TableSet(tw)
    # forward: plain layout [B, T, D] vs the W destination buffers [lN, F, D]
    plain = torch.empty(B, T, D, dtype=dtype, device="cuda")
    ops.emb_fwd(ts, d_idx, d_off, B, plain, D, T * D)
    bufs = [torch.full((lN, F, D), -7.0, dtype=dtype, device="cuda") for _ in range(W)]
    ptrs = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device="cuda")
    for first_dest in range(W):  # the visiting order of the destinations must not change what lands where
        for b_ in bufs:
            b_.fill_(-7.0)
        ops.emb_fwd_sharded(ts, d_idx, d_off, B, ptrs.data_ptr(), lN, F, feat0, dtype, first_dest=first_dest)
        for r in range(W):
            assert torch.equal(bufs[r][:, feat0:feat0 + T], plain[r * lN:(r + 1) * lN])
            assert bool((bufs[r][:, :feat0] == -7).all()) and bool((bufs[r][:, feat0 + T:] == -7).all())
    # backward: gradients scattered over W buffers vs one contiguous gradient
    g = torch.tensor(rs.randint(-3, 4, size=(B, F, D)).astype(np.float32)).to(dtype).cuda()
    gbufs = [g[r * lN:(r + 1) * lN].contiguous() for r in range(W)]
    gptrs = torch.tensor([b.data_ptr() for b in gbufs], dtype=torch.int64, device="cuda")
    tw2 = [w.clone() for w in tw]
    ts2 = ops.TableSet(tw2)
    ws = torch.empty(ops.emb_bwd_workspace_bytes(len(indices), D), dtype=torch.uint8, device="cuda")
    ops.emb_bwd_update(ts, d_idx, d_off, B, g[:, feat0:], D, F * D, lr=0.5, ws=ws)
    ops.emb_bwd_prepare(ts2, d_idx, d_off, B, ws)
    ops.emb_bwd_apply_sharded(ts2, len(indices), B, gptrs.data_ptr(), lN, F, feat0, dtype, 0.5, ws)
    for a, b in zip(tw, tw2):
        assert torch.equal(a, b)


@pytest.mark.parametrize("tag", ["a", "b", "c"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_interaction_itself_and_cat_golden(golden, tag, dtype):
    """--arch-interaction-itself (pairs j <= i) forward / backward and --arch-interaction-op=cat against vectors
    written by the reference's training variant (T/dlrm_s_pytorch.py:407-444, tests/golden/t_options.npz)."""
    ops = _ops()
    from cloudtik_b200.interaction import interaction, interaction_cat
    g = golden("t_options")
    x, ly, R, dR = g[f"{tag}_x"], g[f"{tag}_ly"], g[f"{tag}_R"], g[f"{tag}_dR"]
    rtol, atol = (RTOL32, 1e-5) if dtype == torch.float32 else (2e-2, 0.25)
    xt = torch.tensor(x).to(dtype).cuda().requires_grad_(True)
    lyt = [torch.tensor(t).to(dtype).cuda().requires_grad_(True) for t in ly]
    out = interaction(xt, *lyt, itself=True)
    assert tuple(out.shape) == R.shape
    np.testing.assert_allclose(out.float().detach().cpu().numpy(), R, rtol=rtol, atol=atol)
    # the oracle restatement agrees with the reference too
    np.testing.assert_allclose(ref.interact_fwd(x, list(ly), itself=True), R, rtol=1e-5, atol=1e-5)
    out.backward(torch.tensor(dR).to(dtype).cuda())
    np.testing.assert_allclose(xt.grad.float().cpu().numpy(), g[f"{tag}_dx"], rtol=rtol, atol=atol * 4)
    for k, t in enumerate(lyt):
        np.testing.assert_allclose(t.grad.float().cpu().numpy(), g[f"{tag}_dly"][k], rtol=rtol, atol=atol * 4)
    with torch.no_grad():
        cat = interaction_cat(xt, *lyt)
    np.testing.assert_allclose(cat.float().cpu().numpy(), g[f"{tag}_cat"], rtol=rtol, atol=1e-2 if dtype != torch.float32 else 0)


@pytest.mark.parametrize("B,ragged,inc", [(257, False, False), (300, True, False), (64, True, True)])
def test_linearize_kernel_bit_exact(B, ragged, inc):
    """b200_linearize (one launch) against the oracle's per-table concatenation (P/dlrm_s_pytorch.py:908-915)."""
    ops = _ops()
    rs = np.random.RandomState(B)
    rows = [1000, 37, 5000, 3]
    idx_l, off_l = _make_bags(rs, rows, B, 3, ragged)
    want_i, want_o, want_g = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    base = torch.tensor(np.concatenate([[0], np.cumsum(rows)]), dtype=torch.int64).cuda()
    offs = [torch.tensor(np.concatenate([o, [len(i)]]) if inc else o).cuda() for o, i in zip(off_l, idx_l)]
    gi, go, gg = ops.linearize([torch.tensor(i).cuda() for i in idx_l], offs, [inc] * len(rows), base)
    np.testing.assert_array_equal(gi.cpu().numpy(), want_i)
    np.testing.assert_array_equal(go.cpu().numpy(), want_o)
    np.testing.assert_array_equal(gg.cpu().numpy(), want_g)


def test_survey_named_entry_points_match_update():
    """b200_emb_bwd_sgd / b200_emb_bwd_adagrad / b200_workspace_bytes (the names of SURVEY.md section 8b) give what
    b200_emb_bwd_update gives."""
    ops = _ops()
    from cloudtik_b200._lib import check, lib
    rs = np.random.RandomState(9)
    rows, B, D = [500, 7, 3000], 200, 128
    idx_l, off_l = _make_bags(rs, rows, B, 3, True)
    indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    ti, to = torch.tensor(indices).cuda(), torch.tensor(offsets).cuda()
    g = torch.tensor(rs.randn(B, len(rows), D).astype(np.float32)).cuda()
    assert lib.b200_workspace_bytes(len(indices), D) == ops.emb_bwd_workspace_bytes(len(indices), D)
    ws = torch.empty(ops.emb_bwd_workspace_bytes(len(indices), D), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for opt in ("sgd", "adagrad"):
        W0 = [torch.tensor(rs.uniform(-1, 1, size=(n, D)).astype(np.float32)).cuda() for n in rows]
        S0 = [torch.full((n, D), 0.1, device="cuda") for n in rows]
        a, b = ops.TableSet([w.clone() for w in W0]), ops.TableSet([w.clone() for w in W0])
        if opt == "adagrad":
            a.set_states([s.clone() for s in S0]), b.set_states([s.clone() for s in S0])
        ops.emb_bwd_update(a, ti, to, B, g, D, len(rows) * D, 0.3, ws,
                           optimizer=ops.OPT_ADAGRAD if opt == "adagrad" else ops.OPT_SGD)
        common = (b.d_rows.data_ptr(), b.d_row_base.data_ptr(), ti.data_ptr(), to.data_ptr(), len(indices),
                  g.data_ptr(), len(rows), B, D, D, len(rows) * D, 0, 0)
        if opt == "sgd":
            check(lib.b200_emb_bwd_sgd(b.d_ptrs.data_ptr(), None, *common, 0.3, None, b.total_rows, ws.data_ptr(),
                                       ws.numel(), None, st), "b200_emb_bwd_sgd")
        else:
            check(lib.b200_emb_bwd_adagrad(b.d_ptrs.data_ptr(), b.d_state_ptrs.data_ptr(), None, *common, 0.3, None,
                                           1e-10, b.total_rows, ws.data_ptr(), ws.numel(), None, st),
                  "b200_emb_bwd_adagrad")
        torch.cuda.synchronize()
        for wa, wb in zip(a.weights, b.weights):
            assert torch.equal(wa, wb)


@pytest.mark.parametrize("W", [2, 3, 8])
def test_rowwise_filter_matches_mask_and_prefix_sum(W):
    """b200_rowwise_filter (fixed capacity, fill level on the device) against the torch formulation
    (cloudtik_b200.rowwise.filter_owned: mask + prefix sum); an undersized capacity raises the overflow flag."""
    ops = _ops()
    from cloudtik_b200.rowwise import filter_owned
    rs = np.random.RandomState(W)
    lens = rs.randint(0, 7, size=5000)
    off = torch.tensor(np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)).cuda()
    idx = torch.tensor(rs.randint(0, 100000, size=int(lens.sum())).astype(np.int64)).cuda()
    for r in range(W):
        want_i, want_o = filter_owned(idx, off, W, r)
        err = torch.zeros(1, dtype=torch.int32, device="cuda")
        li, lo = ops.rowwise_filter(idx, off, W, r, idx.numel(), err)
        n = int(lo[-1].item())
        assert n == want_i.numel() and int(err.item()) == 0
        assert torch.equal(lo, want_o) and torch.equal(li[:n], want_i)
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    li, lo = ops.rowwise_filter(idx, off, W, 0, 10, err)
    assert int(err.item()) == 2 and int(lo[-1].item()) == 10


@pytest.mark.parametrize("opt", ["sgd", "adagrad"])
def test_partial_prepare_stops_at_the_fill_level(opt):
    """A fixed-capacity index array filled up to a device-side count (b200_emb_bwd_prepare_partial): the update must
    equal the one of the exactly sized array -- the tail sorts behind the real keys and is never touched."""
    ops = _ops()
    rs = np.random.RandomState(3)
    rows, B, D = [700, 5, 3000], 300, 128
    idx_l, off_l = _make_bags(rs, rows, B, 4, True)
    indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    n = len(indices)
    g = torch.tensor(rs.randn(B, len(rows), D).astype(np.float32)).cuda()
    W0 = [torch.tensor(rs.uniform(-1, 1, size=(r, D)).astype(np.float32)).cuda() for r in rows]
    res = []
    for cap in (n, n + 777):
        ts = ops.TableSet([w.clone() for w in W0])
        if opt == "adagrad":
            ts.set_states([torch.full((r, D), 0.1, device="cuda") for r in rows])
        ind = torch.full((cap,), 123456789, dtype=torch.int64, device="cuda")  # garbage behind the fill level
        ind[:n] = torch.tensor(indices)
        off = torch.tensor(offsets).cuda()
        ws = torch.empty(ops.emb_bwd_workspace_bytes(cap, D), dtype=torch.uint8, device="cuda")
        ops.emb_bwd_prepare(ts, ind, off, B, ws, nnz_valid=off[-1:] if cap != n else None)
        ops.emb_bwd_apply(ts, cap, B, g, D, len(rows) * D, 0.2, ws,
                          optimizer=ops.OPT_ADAGRAD if opt == "adagrad" else ops.OPT_SGD)
        torch.cuda.synchronize()
        res.append([w.clone() for w in ts.weights])
    for a_, b_ in zip(*res):
        assert torch.equal(a_, b_)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_rowwise_partial_sums_add_up_over_emulated_ranks(dtype):
    """b200_emb_fwd_sharded_add on one GPU with W emulated ranks: every "rank" filters the global bags down to its
    rows and ADDS its partial pooled rows into the destination buffers; together they must give the pooled rows of
    the unsharded table (the reduce-scatter of the row-wise scheme)."""
    ops = _ops()
    W, lN, D, k = 4, 32, 128, 2
    F = 1 + k
    B = W * lN
    rs = np.random.RandomState(12)
    rows = [1000, 37]
    full = [rs.uniform(-1, 1, size=(n, D)).astype(np.float32) for n in rows]
    idx_l, off_l = _make_bags(rs, rows, B, 5, True)
    # global row ids per table are table-local here: one filter call per table keeps ids table-local
    T_dst = [torch.zeros(lN, F, D, dtype=dtype, device="cuda") for _ in range(W)]
    ptrs = torch.tensor([t.data_ptr() for t in T_dst], dtype=torch.int64, device="cuda")
    for r in range(W):
        shards = [torch.tensor(np.ascontiguousarray(f[r::W])).cuda() for f in full]
        ts = ops.TableSet(shards)
        lis, los, base = [], [], 0
        for t in range(k):
            off = torch.tensor(np.concatenate([off_l[t], [len(idx_l[t])]]).astype(np.int64)).cuda()
            li, lo = ops.rowwise_filter(torch.tensor(idx_l[t]).cuda(), off, W, r, len(idx_l[t]) + 2)
            n = int(lo[-1].item())
            lis.append(li[:n])
            los.append(lo[:-1] + base)
            base += n
        li = torch.cat(lis)
        lo = torch.cat(los + [torch.tensor([base], dtype=torch.int64, device="cuda")])
        ops.emb_fwd_sharded(ts, li, lo, B, ptrs.data_ptr(), lN, F, 1, dtype, accumulate=True)
    torch.cuda.synchronize()
    got = torch.cat(T_dst, dim=0).float().cpu().numpy()  # [B, F, D]
    for t in range(k):
        want = ref.embedding_bag_sum(full[t], idx_l[t], off_l[t])
        tol = dict(rtol=1e-5, atol=1e-5) if dtype == torch.float32 else dict(rtol=3e-2, atol=6e-2)
        np.testing.assert_allclose(got[:, 1 + t], want, **tol)
    assert not got[:, 0].any()


@pytest.mark.parametrize("variant", ["sgd_fp32", "adagrad_bf16grad", "bf16_trail", "dense"])
def test_emb_bwd_very_hot_rows_cta_join(variant):
    """Rows with tens of thousands of duplicates (multi-hot bags on a 3-row table: runs that cross far more than
    HOT_BLOCKS block borders) take the CTA-wide hot join; integer gradients make the sums exact."""
    ops = _ops()
    D, rows, B, P = 128, [3, 50, 7], 6000, 8
    rs = np.random.RandomState(1)
    idx_l = [rs.randint(0, n, size=B * P).astype(np.int64) for n in rows]
    off_l = [np.arange(B, dtype=np.int64) * P for _ in rows]
    indices, offsets, _ = ref.linearize_indices_and_offsets(idx_l, off_l, rows)
    T = len(rows)
    gdt = torch.bfloat16 if variant in ("adagrad_bf16grad", "bf16_trail") else torch.float32
    g = torch.tensor(rs.randint(-2, 3, size=(B, T, D)).astype(np.float32)).to(gdt).cuda()
    gr = g.float().cpu().numpy().astype(np.float64)
    G = []
    for t in range(T):
        Gt = np.zeros((rows[t], D))
        np.add.at(Gt, idx_l[t], np.repeat(gr[:, t], P, axis=0))
        G.append(Gt)
    ti, to = torch.tensor(indices).cuda(), torch.tensor(offsets).cuda()
    ws = torch.empty(ops.emb_bwd_workspace_bytes(len(indices), D), dtype=torch.uint8, device="cuda")
    if variant == "dense":
        ts = ops.TableSet([torch.zeros(n, D, device="cuda") for n in rows])
        ops.emb_bwd_prepare(ts, ti, to, B, ws)
        out = torch.zeros(sum(rows), D, device="cuda")
        ops.emb_bwd_dense_grad(ts, len(indices), B, g, D, T * D, out, ws)
        np.testing.assert_array_equal(out.cpu().numpy(), np.concatenate(G))
        return
    lr = 2.0 ** -10
    if variant == "bf16_trail":
        W = [torch.zeros(n, D, dtype=torch.bfloat16, device="cuda") for n in rows]
        ts = ops.TableSet(W)
        trails = [torch.zeros(n, D, dtype=torch.int16, device="cuda") for n in rows]
        ts.set_trails(trails)
    else:
        W = [torch.zeros(n, D, device="cuda") for n in rows]
        ts = ops.TableSet(W)
    opt = ops.OPT_SGD
    if variant == "adagrad_bf16grad":
        ts.set_states([torch.zeros(n, D, device="cuda") for n in rows])
        opt = ops.OPT_ADAGRAD
    ops.emb_bwd_update(ts, ti, to, B, g, D, T * D, lr, ws, optimizer=opt)
    torch.cuda.synchronize()
    for t in range(T):
        if variant == "adagrad_bf16grad":
            np.testing.assert_array_equal(ts.states[t].cpu().numpy(), (G[t] * G[t]).astype(np.float32))
            want = -lr * G[t] / (np.abs(G[t]) + 1e-10)
            np.testing.assert_allclose(W[t].cpu().numpy(), want, rtol=1e-5, atol=1e-7)
        elif variant == "bf16_trail":
            hi = W[t].view(torch.int16).to(torch.int32) << 16
            lo = trails[t].to(torch.int32) & 0xFFFF
            np.testing.assert_array_equal((hi | lo).view(torch.float32).cpu().numpy(), (-lr * G[t]).astype(np.float32))
        else:
            np.testing.assert_array_equal(W[t].cpu().numpy(), (-lr * G[t]).astype(np.float32))


@pytest.mark.parametrize("B,F,D,pad", [(1, 27, 128, 1), (5, 27, 128, 8), (1000, 27, 128, 8), (4099, 27, 128, 1),
                                       (333, 27, 64, 8), (777, 9, 32, 1), (64, 32, 128, 1)])
def test_interaction_fp32_on_tensor_cores_3xtf32(B, F, D, pad):
    """fp32 forward on tcgen05 (3xTF32, b200_interact_fwd_tf32) against a float64 evaluation of the same inputs at
    the north_star fp32 bound (1e-5 relative) -- the same check as the CUDA-core fp32 kernel."""
    from cloudtik_b200._lib import check, lib
    rs = np.random.RandomState(B + D)
    Tn = rs.randn(B, F, D).astype(np.float32)
    T = torch.tensor(Tn).cuda()
    W = D + F * (F - 1) // 2
    ldr = (W + pad - 1) // pad * pad
    R = torch.full((B, ldr), 7.0, device="cuda")
    check(lib.b200_interact_fwd_tf32(T.data_ptr(), R.data_ptr(), B, F, D, ldr, torch.cuda.current_stream().cuda_stream),
          "b200_interact_fwd_tf32")
    T64 = Tn.astype(np.float64)
    want = ref.interact_fwd(T64[:, 0], [T64[:, k] for k in range(1, F)])
    got = R.cpu().numpy()
    np.testing.assert_allclose(got[:, :W], want, rtol=1e-5, atol=1e-5 * np.sqrt(D))
    np.testing.assert_array_equal(got[:, :D], Tn[:, 0])
    assert (got[:, W:] == 0).all()
