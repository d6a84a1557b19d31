"""The reference's own test configurations (T/test/dlrm_s_test.sh:24,26) through the B200 model: MSE loss, D = 2,
ragged bags of up to 10 indices, batches of 2 and 1 samples -- losses the reference printed, within 1e-5."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag", ["c2", "c4"])
def test_reference_test_script_configs(golden, tag):
    from cloudtik_b200.dlrm import DLRM_Net, train_step
    g = golden("t_smoke")
    np.random.seed(0)
    net = DLRM_Net(2, g["ln_emb"], g["ln_bot"], g["ln_top"], sigmoid_top=len(g["ln_top"]) - 2,
                   np_init_emb_weight=True, device="cuda", lr=0.1)
    init = {str(n): torch.tensor(g[f"{tag}_init{k}"]) for k, n in enumerate(g[f"{tag}_param_names"])}
    net.load_state_dict(init)
    net.loss_fn = torch.nn.MSELoss(reduction="mean")
    opt = torch.optim.SGD([p for p in net.parameters() if p.requires_grad], lr=0.1)
    losses = []
    for s in range(int(g[f"{tag}_steps"])):
        lS_o = [torch.tensor(g[f"{tag}_o{s}_{t}"]).cuda() for t in range(3)]
        lS_i = [torch.tensor(g[f"{tag}_i{s}_{t}"]).cuda() for t in range(3)]
        losses.append(float(train_step(net, opt, torch.tensor(g[f"{tag}_X{s}"]).cuda(), (lS_o, lS_i),
                                       torch.tensor(g[f"{tag}_T{s}"]).cuda()).cpu()))
    np.testing.assert_allclose(np.array(losses), g[f"{tag}_losses"], rtol=1e-5, atol=1e-7)
    sd = net.state_dict()
    for k, n in enumerate(g[f"{tag}_param_names"]):
        np.testing.assert_allclose(sd[str(n)].float().cpu().numpy(), g[f"{tag}_final{k}"], rtol=1e-5, atol=1e-6)
