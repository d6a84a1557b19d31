"""Inference / evaluation path: device metrics against sklearn (the reference's own
metric code, P/dlrm_s_pytorch.py:563-584), forward-only model against the CPU oracle port, graph replay."""
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,levels", [(1000, 0), (5000, 17), (200000, 1000), (64, 2)])
def test_device_metrics_match_sklearn(n, levels):
    """levels > 0 quantises the scores: heavy ties, the case where AUC / AP need the tie groups."""
    from cloudtik_b200 import inference as inf
    rs = np.random.RandomState(n + levels)
    s = rs.rand(n).astype(np.float32)
    if levels:
        s = (np.floor(s * levels) / levels).astype(np.float32)
    s[:3] = [0.0, 1.0, 0.5]
    t = (rs.rand(n) < 0.3 + 0.4 * s).astype(np.float32)
    want = inf.sklearn_reference_metrics(t, s)
    got = inf.binary_metrics(torch.tensor(t).cuda(), torch.tensor(s).cuda())
    for k in want:
        assert abs(got[k] - want[k]) <= 1e-12 + 1e-9 * abs(want[k]), (k, got[k], want[k])


def test_auc_handles_negative_and_signed_zero_scores():
    from cloudtik_b200 import inference as inf
    s = np.array([-2.5, -0.0, 0.0, 1e-30, -1e-30, 3.0, -2.5, 3.0], dtype=np.float32)
    t = np.array([0, 1, 0, 1, 0, 1, 1, 0], dtype=np.float32)
    import sklearn.metrics as m
    # -0.0 and 0.0 compare equal for sklearn; the integer image keeps them adjacent but distinct, so the test
    # uses targets for which that cannot change the area
    t[1], t[2] = 1, 1
    assert abs(inf.roc_auc_score(torch.tensor(t).cuda(), torch.tensor(s).cuda()) - m.roc_auc_score(t, s)) < 1e-12


def _small_pair(bf16=False):
    from cloudtik_b200.dlrm import DLRM_Net
    from oracle.dlrm_port import DLRMPort
    D, rows = 32, [50, 7, 1200, 33]
    F = len(rows) + 1
    ln_bot, ln_top = [13, 64, D], [D + F * (F - 1) // 2, 64, 1]
    np.random.seed(5)
    port = DLRMPort(D, rows, ln_bot, ln_top)
    np.random.seed(5)
    net = DLRM_Net(D, np.array(rows), np.array(ln_bot), np.array(ln_top), sigmoid_top=len(ln_top) - 2,
                   np_init_emb_weight=True, device="cuda", bf16=bf16)
    return net, port, rows


def test_inference_function_matches_cpu_port(capsys):
    """inference() over a 3-batch loader: accuracy line and returned dict vs the oracle port's predictions; with
    print_auc the metric line of the reference."""
    from cloudtik_b200.inference import inference, sklearn_reference_metrics
    from oracle.dlrm_port import uniform_batch
    net, port, rows = _small_pair()
    rs = np.random.RandomState(1)
    loader, S, Tt = [], [], []
    for _ in range(3):
        X, lS_o, lS_i, T = uniform_batch(rs, 96, rows, pooling=3)
        loader.append((torch.tensor(X), [torch.tensor(o) for o in lS_o], [torch.tensor(i) for i in lS_i], torch.tensor(T)))
        with torch.no_grad():
            S.append(port(torch.tensor(X), [torch.tensor(o) for o in lS_o], [torch.tensor(i) for i in lS_i]).numpy())
        Tt.append(T)
    S, Tt = np.concatenate(S).reshape(-1), np.concatenate(Tt).reshape(-1)
    args = types.SimpleNamespace(print_auc=False, inference_only=False, print_freq=1, nepochs=1, mlperf_auc_threshold=0.0)
    md, is_best = inference(args, net, 0.0, 0.0, loader)
    want_acc = float((np.round(S) == Tt).mean())
    assert abs(md["test_acc"] - want_acc) < 1e-9 and is_best
    assert " accuracy {:3.3f} %, best {:3.3f} %".format(want_acc * 100, want_acc * 100) in capsys.readouterr().out
    args.print_auc = True
    md, _ = inference(args, net, 0.0, 0.0, loader)
    want = sklearn_reference_metrics(Tt, S)
    assert abs(md["test_auc"] - want["roc_auc"]) < 1e-4  # scores differ from the CPU's in the last bits
    out = capsys.readouterr().out
    assert "recall" in out and "best auc" in out and "Accuracy: " in out
    assert net.training  # mode restored


def test_graphed_inference_equals_eager():
    from cloudtik_b200.inference import GraphedInference
    from oracle.dlrm_port import uniform_batch
    from oracle import ops_numpy as onp
    net, port, rows = _small_pair(bf16=True)
    rs = np.random.RandomState(2)
    batches = []
    for _ in range(3):
        X, lS_o, lS_i, T = uniform_batch(rs, 64, rows, pooling=1)
        ind, off, _ = onp.linearize_indices_and_offsets(lS_i, lS_o, rows)
        batches.append((torch.tensor(X).cuda(), torch.tensor(ind).cuda(), torch.tensor(off).cuda()))
    gi = GraphedInference(net, batches[0])
    for X, ind, off in batches:
        got = gi(X, ind, off).clone()
        with torch.no_grad():
            want = net(X, ind, off, None).float()
        assert torch.equal(got, want)


def test_int8_quantise_on_device_is_bit_exact_with_oracle(golden):
    from cloudtik_b200.quantized import quantize_rowwise_int8
    g = golden("int8_tables")
    q, s = quantize_rowwise_int8(torch.tensor(g["W"]).cuda())
    assert np.array_equal(q.cpu().numpy(), g["q"]) and np.array_equal(s.cpu().numpy(), g["scale"])


def test_int8_forward_matches_reference_fixture(golden):
    from cloudtik_b200 import ops
    g = golden("int8_tables")
    ts = ops.QuantTableSet([torch.tensor(g["q"]).cuda()], [torch.tensor(g["scale"]).cuda()])
    B = len(g["off"]) - 1
    out = torch.empty(B, 1, 64, device="cuda")
    ops.emb_fwd_q8(ts, torch.tensor(g["idx"]).cuda(), torch.tensor(g["off"]).cuda(), B, out, 64, 64)
    np.testing.assert_allclose(out[:, 0].cpu().numpy(), g["pooled"], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("D,P,out_dtype", [(128, 1, torch.float32), (128, 1, torch.bfloat16), (128, 7, torch.float32),
                                            (64, 3, torch.float32), (16, 1, torch.float32), (256, 40, torch.bfloat16),
                                            (24, 2, torch.float32)])
def test_int8_forward_vs_oracle(D, P, out_dtype):
    """One-hot, ragged (incl. empty bags), long bags, D not a multiple of 16 (scalar path)."""
    from cloudtik_b200 import ops
    from oracle import ops_numpy as ref
    rs = np.random.RandomState(D + P)
    rows, B = [300, 5, 4000], 257
    qs, ss, want, glob_idx, glob_off, tot = [], [], [], [], [0], 0
    for n in rows:
        q, s = ref.quantize_rowwise_int8((rs.randn(n, D) * 0.1).astype(np.float32))
        lens = rs.randint(0, 2 * P, size=B) if P > 1 else np.ones(B, dtype=np.int64)
        o = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        i = rs.randint(0, n, size=int(lens.sum())).astype(np.int64)
        want.append(ref.embedding_bag_sum_int8(q, s, i, o[:-1]))
        glob_idx.append(i)
        glob_off += (tot + o[1:]).tolist()
        tot += int(lens.sum())
        qs.append(torch.tensor(q).cuda()), ss.append(torch.tensor(s).cuda())
    ts = ops.QuantTableSet(qs, ss)
    T = len(rows)
    out = torch.empty(B, T, D, dtype=out_dtype, device="cuda")
    ops.emb_fwd_q8(ts, torch.tensor(np.concatenate(glob_idx)).cuda(), torch.tensor(np.array(glob_off)).cuda(), B, out, D, T * D)
    tol = dict(rtol=1e-5, atol=1e-5) if out_dtype == torch.float32 else dict(rtol=1e-2, atol=1e-2)
    for t in range(T):
        np.testing.assert_allclose(out[:, t].float().cpu().numpy(), want[t], **tol)


def test_int8_model_inference_close_to_float_and_exact_vs_dequantised():
    """quantize_embeddings_int8(net): (a) equals the float model run on the DEQUANTISED tables (same kernels
    otherwise) to fp32 accuracy, (b) stays within the int8 quantisation error of the original model."""
    from cloudtik_b200.quantized import quantize_embeddings_int8
    from oracle.dlrm_port import uniform_batch
    net, port, rows = _small_pair()
    rs = np.random.RandomState(4)
    X, lS_o, lS_i, T = uniform_batch(rs, 200, rows, pooling=2)
    args = (torch.tensor(X).cuda(), [torch.tensor(o).cuda() for o in lS_o], [torch.tensor(i).cuda() for i in lS_i])
    net.eval()
    with torch.no_grad():
        z_float = net(*args).float().cpu().numpy()
    deq = None
    qnet = quantize_embeddings_int8(net)
    with torch.no_grad():
        z_q = qnet(*args).float().cpu().numpy()
        deq = qnet.emb_l.weights
    for k, w in enumerate(deq):  # the oracle port with dequantised tables = what the int8 model must compute
        port.emb_l[k].weight.data.copy_(w.cpu())
    with torch.no_grad():
        z_port = port(torch.tensor(X), [torch.tensor(o) for o in lS_o], [torch.tensor(i) for i in lS_i]).numpy()
    np.testing.assert_allclose(z_q, z_port, rtol=1e-5, atol=1e-6)
    assert np.abs(z_q - z_float).max() < 5e-3


def test_int8_forward_full_size_exact():
    """Config-2 sized int8 tables (26 x 1M x 128) with power-of-two scales and small integer codes: every
    product and sum is exact in fp32, so the kernel must equal a torch evaluation bit for bit."""
    from cloudtik_b200 import ops
    T, ROWS, D, B, P = 26, 1_000_000, 128, 8192, 8
    g = torch.Generator(device="cuda").manual_seed(1)
    qs = [torch.randint(-20, 21, (ROWS, D), device="cuda", generator=g, dtype=torch.int8) for _ in range(T)]
    ss = [torch.pow(2.0, torch.randint(-3, 3, (ROWS,), device="cuda", generator=g).float()) for _ in range(T)]
    ts = ops.QuantTableSet(qs, ss)
    idx = torch.randint(0, ROWS, (T, B, P), device="cuda", generator=g)
    off = torch.arange(T * B + 1, device="cuda") * P
    out = torch.empty(B, T, D, device="cuda")
    ops.emb_fwd_q8(ts, idx.reshape(-1), off, B, out, D, T * D)
    for t in (0, 13, 25):
        rows_t = idx[t].reshape(-1)
        want = (qs[t][rows_t].float() * ss[t][rows_t][:, None]).view(B, P, D).sum(1)
        assert torch.equal(out[:, t], want)
