"""Host-side logic of cloudtik_b200.mlp that needs no GPU: the bf16 parameter shadows are used only while fresh."""
import torch
import torch.nn as nn


def test_shadow_freshness_and_padding():
    from cloudtik_b200.interaction import PaddedLinear
    from cloudtik_b200.mlp import _bf16_of, enable_bf16_shadows
    torch.manual_seed(0)
    mlp = nn.Sequential(PaddedLinear(13, 8), nn.ReLU(), nn.Linear(8, 4), nn.Sigmoid())
    opt = torch.optim.SGD(mlp.parameters(), lr=0.1)
    enable_bf16_shadows([mlp], opt)
    w0, w1 = mlp[0].weight, mlp[2].weight
    sh, fresh = _bf16_of(w0, 16)                       # 13 -> 16 columns, zero padded
    assert fresh and sh.shape == (8, 16) and torch.equal(sh[:, :13], w0.detach().bfloat16()) and not sh[:, 13:].any()
    assert not _bf16_of(w0, 13)[1]                     # unpadded input: the padded shadow does not fit -> cast
    sh1, fresh1 = _bf16_of(w1, 8)
    assert fresh1 and torch.equal(sh1, w1.detach().bfloat16())
    # a training step refreshes every shadow through the optimizer's post-hook
    mlp(torch.randn(5, 13)).sum().backward()
    opt.step()
    assert _bf16_of(w1, 8)[1] and torch.equal(_bf16_of(w1, 8)[0], w1.detach().bfloat16())
    assert torch.equal(_bf16_of(mlp[2].bias)[0], mlp[2].bias.detach().bfloat16())
    # a change behind the optimizer's back makes the shadow stale: the helper falls back to a cast
    with torch.no_grad():
        w1.mul_(2.0)
    val, fresh = _bf16_of(w1, 8)
    assert not fresh and torch.equal(val, w1.detach().bfloat16())
    opt.step()                                         # next step: fresh again
    assert _bf16_of(w1, 8)[1]
