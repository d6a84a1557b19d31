"""QR / mixed-dimension embedding operators against fixtures produced by the reference's
own T/tricks modules on the CPU: forward, sparse gradients of both tables (indices bit-exact, values 1e-5)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("op", ["mult", "add", "concat"])
def test_qr_embedding_bag_matches_reference(golden, op):
    from cloudtik_b200.tricks import QREmbeddingBag
    g = golden("tricks")
    ncat, C, D = int(g[f"qr_{op}_ncat"]), 7, 16
    nq = int(np.ceil(ncat / C))
    wq = torch.full((nq, D), 0.25)
    wq[torch.tensor(g[f"qr_{op}_q_rows"])] = torch.tensor(g[f"qr_{op}_wq_rows"])
    m = QREmbeddingBag(ncat, D, C, operation=op, mode="sum", sparse=True, _weight=[wq, torch.tensor(g[f"qr_{op}_wr"])])
    y = m(torch.tensor(g[f"qr_{op}_idx"]).cuda(), torch.tensor(g[f"qr_{op}_off"]).cuda())
    np.testing.assert_allclose(y.detach().cpu().numpy(), g[f"qr_{op}_y"], rtol=1e-5, atol=1e-5)
    y.backward(torch.tensor(g[f"qr_{op}_gy"]).cuda())
    for name, w in (("q", m.weight_q), ("r", m.weight_r)):
        gr = w.grad
        # ids above 2**24 go through the reference's float32 quotient: same (inexact) rows here
        np.testing.assert_array_equal(gr._indices().cpu().numpy()[0], g[f"qr_{op}_g{name}_idx"])
        np.testing.assert_allclose(gr._values().cpu().numpy(), g[f"qr_{op}_g{name}_val"], rtol=1e-5, atol=1e-5)


def test_pr_embedding_bag_matches_reference(golden):
    from cloudtik_b200.tricks import PrEmbeddingBag
    g = golden("tricks")
    m = PrEmbeddingBag(500, 8, 32)
    with torch.no_grad():
        m.embs.weight.copy_(torch.tensor(g["pr_w"]))
        m.proj.weight.copy_(torch.tensor(g["pr_proj"]))
    y = m(torch.tensor(g["pr_idx"]).cuda(), torch.tensor(g["pr_off"]).cuda())
    np.testing.assert_allclose(y.detach().cpu().numpy(), g["pr_y"], rtol=1e-5, atol=1e-5)
    y.backward(torch.tensor(g["pr_gy"]).cuda())
    np.testing.assert_array_equal(m.embs.weight.grad._indices().cpu().numpy()[0], g["pr_ge_idx"])
    np.testing.assert_allclose(m.embs.weight.grad._values().cpu().numpy(), g["pr_ge_val"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(m.proj.weight.grad.cpu().numpy(), g["pr_gproj"], rtol=1e-5, atol=1e-5)
    assert isinstance(PrEmbeddingBag(10, 32, 32).proj, torch.nn.Identity)
    with pytest.raises(ValueError):
        PrEmbeddingBag(10, 64, 32)
