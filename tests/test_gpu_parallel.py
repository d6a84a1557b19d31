"""Single-process multi-GPU mode (``parallel_forward`` T/dlrm_s_pytorch.py:524-630):
the model spread over two GPUs by ONE process must train like the same model on one GPU.  Needs two visible GPUs
(skipped on a one-GPU box); run with ``gpurun --gpus 2 -- python -m pytest tests/test_gpu_parallel.py``."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

two_gpus = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")


def _batch(rs, ln_emb, B, P, m_den):
    X = torch.tensor(rs.rand(B, m_den).astype(np.float32))
    lS_o, lS_i = [], []
    for rows in ln_emb:
        sizes = rs.randint(1, P + 1, size=B)
        off = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64)
        idx = rs.randint(0, rows, size=int(sizes.sum())).astype(np.int64)
        lS_o.append(torch.tensor(off))
        lS_i.append(torch.tensor(idx))
    T = torch.tensor(np.round(rs.rand(B, 1)).astype(np.float32))
    return X, lS_o, lS_i, T


def _tables(net):
    """Tables of a model in natural order, after every pending update."""
    if net.ndevices > 1:
        out = {}
        for d, ks in enumerate(net.tables_of_device):
            net.emb_l[d].join()
            for pos, k in enumerate(ks):
                out[k] = net.emb_l[d].weights[pos].detach().float().cpu()
        for d in net.device_ids:
            torch.cuda.synchronize(d)
        return [out[k] for k in sorted(out)]
    net.emb_l.join()
    torch.cuda.synchronize()
    return [w.detach().float().cpu() for w in net.emb_l.weights]


@two_gpus
@pytest.mark.parametrize("bf16,D,B", [(False, 32, 51), (True, 128, 64), (False, 128, 7)])
def test_parallel_forward_trains_like_one_gpu(bf16, D, B):
    from cloudtik_b200.dlrm import DLRM_Net
    torch.cuda.set_device(0)
    ln_emb = np.array([37, 1000, 5, 600, 123])
    F = 1 + len(ln_emb)
    ln_bot = np.array([13, 64, D])
    ln_top = np.array([D + F * (F - 1) // 2, 64, 32, 1])

    def build(nd):
        np.random.seed(11)
        torch.manual_seed(11)
        return DLRM_Net(D, ln_emb, ln_bot, ln_top, sigmoid_top=len(ln_top) - 2, np_init_emb_weight=True,
                        device="cuda", lr=0.1, bf16=bf16, ndevices=nd)

    one, two = build(-1), build(2)
    assert two.ndevices == 2 and two.tables_of_device == [[0, 2, 4], [1, 3]]
    assert [w.device.index for w in two.emb_l[1].weights] == [1, 1]
    for a, b in zip(_tables(one), _tables(two)):
        np.testing.assert_array_equal(a.numpy(), b.numpy())  # same seed, same tables (natural draw order)
    opts = [torch.optim.SGD([p for p in n.parameters() if p.requires_grad], lr=0.1) for n in (one, two)]
    rs = np.random.RandomState(5)
    tol = 3e-2 if bf16 else 1e-5
    for _step in range(3):
        X, lS_o, lS_i, T = _batch(rs, ln_emb, B, 4, 13)
        losses = []
        for net, opt in zip((one, two), opts):
            Z = net(X.cuda(0), [o.cuda(0) for o in lS_o], [i.cuda(0) for i in lS_i])
            assert Z.device.index == 0 and tuple(Z.shape) == (B, 1)
            E = net.loss_fn(Z.float(), T.cuda(0))
            opt.zero_grad()
            E.backward()
            opt.step()
            losses.append(float(E.detach()))
        assert abs(losses[0] - losses[1]) <= tol * abs(losses[0]), losses
    for a, b in zip(_tables(one), _tables(two)):
        torch.testing.assert_close(b, a, rtol=tol, atol=5e-3 if bf16 else 1e-6)
    for pa, pb in zip(one.parameters(), two.parameters()):
        if pa.requires_grad:
            torch.testing.assert_close(pb.detach().cpu(), pa.detach().cpu(), rtol=10 * tol, atol=5e-3 if bf16 else 5e-6)
    # forward only
    X, lS_o, lS_i, _ = _batch(rs, ln_emb, B, 4, 13)
    with torch.no_grad():
        za = one(X.cuda(0), [o.cuda(0) for o in lS_o], [i.cuda(0) for i in lS_i]).float().cpu()
        zb = two(X.cuda(0), [o.cuda(0) for o in lS_o], [i.cuda(0) for i in lS_i]).float().cpu()
    torch.testing.assert_close(zb, za, rtol=tol, atol=2e-2 if bf16 else 1e-6)


@two_gpus
def test_parallel_forward_refuses_what_the_reference_refuses():
    from cloudtik_b200.dlrm import DLRM_Net
    torch.cuda.set_device(0)
    ln_emb = np.array([10, 20, 30])
    np.random.seed(1)
    net = DLRM_Net(16, ln_emb, np.array([4, 16]), np.array([16 + 6, 8, 1]), sigmoid_top=1, np_init_emb_weight=True,
                   device="cuda", ndevices=2)
    X, lS_o, lS_i, _ = _batch(np.random.RandomState(0), ln_emb, 4, 2, 4)
    with pytest.raises(SystemExit):  # T/dlrm_s_pytorch.py:556-557
        net(X.cuda(0), [o.cuda(0) for o in lS_o[:2]], [i.cuda(0) for i in lS_i[:2]])
    with pytest.raises(SystemExit):
        net(X[:1].cuda(0), [o[:1].cuda(0) for o in lS_o], [i[:1].cuda(0) for i in lS_i])
