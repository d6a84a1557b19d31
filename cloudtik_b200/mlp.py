"""bf16 fast path for the dense MLPs (they stay on torch / cuBLAS, north_star).

The reference builds ``nn.Sequential(Linear, ReLU, ..., Linear, Sigmoid)`` (``P/dlrm_s_pytorch.py:172-211``)
and runs it under autocast.  On a B200 the GEMMs are a minor part of that: per layer autocast also
launches a ReLU kernel, two cast/copy kernels and -- the expensive one -- a column reduction of the
bf16 ``[B, N]`` output gradient for the bias gradient (48 us for 32768 x 1024, as much as the GEMM).

``mlp_forward`` walks the SAME modules (same parameters, same state_dict keys) but calls, per
(Linear, ReLU) pair, one autograd function that
  * fuses bias + ReLU into the GEMM epilogue (``torch._addmm_activation``),
  * masks the output gradient with the ReLU and sums its columns for the bias gradient in ONE pass over the
    matrix (``b200_relu_bwd_bias``; layers without a ReLU use a skinny tensor-core GEMM ``ones[8, B] @ dY``),
  * casts each weight to bf16 once per step.
Numerically it is the autocast path (bf16 operands, fp32 accumulation, bf16 activations / gradients).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops
from .interaction import PaddedLinear

_ONES = {}

# ---- bf16 shadows of the fp32 MLP parameters ------------------------------------------------------
# Autocast-style training casts every weight and bias to bf16 in every forward: 16 tiny kernels per step for
# the 8 DLRM layers (plus a zero-pad for the two K-padded ones).  enable_bf16_shadows() keeps a bf16 copy of
# each parameter (already padded), refreshed by ONE multi-tensor copy right after optimizer.step() (the
# reference's IPEX path keeps bf16 weights next to fp32-precision masters the same way, P/dlrm_s_pytorch.py:
# 926-953).  A shadow is used only while its recorded version matches the parameter's: any other in-place
# change of a parameter (load_state_dict, manual edits) silently falls back to the per-step cast.
_SHADOWS = {}  # id(param) -> [weakref(param), shadow, version, pad]


def enable_bf16_shadows(layers_list, optimizer):
    """``layers_list``: the model's nn.Sequential MLPs; ``optimizer``: the torch optimizer that steps their
    parameters.  Returns the refresh function (also installed as a step post-hook of ``optimizer``)."""
    import weakref
    entries = []
    for layers in layers_list:
        mods = list(layers)
        for i, m in enumerate(mods):
            if not isinstance(m, (nn.Linear, PaddedLinear)):
                continue
            pad = (-m.in_features) % 8 if isinstance(m, PaddedLinear) else 0
            w = m.weight
            sh = torch.zeros(w.shape[0], w.shape[1] + pad, dtype=torch.bfloat16, device=w.device)
            entries.append([weakref.ref(w), sh, -1, pad])
            _SHADOWS[id(w)] = entries[-1]
            if m.bias is not None:
                entries.append([weakref.ref(m.bias), torch.zeros_like(m.bias, dtype=torch.bfloat16), -1, 0])
                _SHADOWS[id(m.bias)] = entries[-1]

    def refresh(*_a, **_k):
        dst, src = [], []
        for e in entries:
            p = e[0]()
            if p is None:
                continue
            if e[3]:
                e[1][:, :p.shape[1]].copy_(p.data)  # padded weight: strided destination, its own copy
            else:
                dst.append(e[1]), src.append(p.data)
            e[2] = p._version
        if dst:
            torch._foreach_copy_(dst, src)
    refresh()
    optimizer.register_step_post_hook(refresh)
    return refresh


def _bf16_of(param, width=None):
    """bf16 value of a parameter: its shadow when fresh (and of the requested padded width), else a cast."""
    e = _SHADOWS.get(id(param))
    if e is not None and e[0]() is param and e[2] == param._version and (width is None or e[1].shape[1] == width):
        return e[1], True
    return param.to(torch.bfloat16), False



def _ones(rows: int, device) -> torch.Tensor:
    key = (rows, device)
    t = _ONES.get(key)
    if t is None:
        t = _ONES[key] = torch.ones(8, rows, dtype=torch.bfloat16, device=device)
    return t


class _LinearReLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, relu: bool):
        x16 = x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16)
        extra = x16.shape[1] - weight.shape[1]
        w16, padded = _bf16_of(weight, x16.shape[1])
        if extra > 0 and not padded:  # PaddedLinear: input carries zero columns beyond in_features
            w16 = torch.nn.functional.pad(w16, (0, extra))
        b16, _ = _bf16_of(bias)
        if relu:
            y = torch._addmm_activation(b16, x16, w16.t())  # bias + ReLU in the GEMM epilogue
        else:
            y = torch.addmm(b16, x16, w16.t())
        ctx.save_for_backward(x16, w16, y)
        ctx.relu = relu
        ctx.extra = extra
        ctx.x_dtype = x.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        x16, w16, y = ctx.saved_tensors
        dy = dy.contiguous()
        dx = dw = db = None
        if ctx.relu:
            if dy.dtype == torch.bfloat16 and dy.shape[1] % 8 == 0 and dy.data_ptr() % 16 == 0:
                # one pass: ReLU mask + fp32 column sums (b200_relu_bwd_bias) instead of threshold_backward
                # followed by a bias-gradient reduction of the same matrix
                dy, db = ops.relu_bwd_bias(dy, y)
            else:
                dy = torch.ops.aten.threshold_backward(dy, y, 0)
        if ctx.needs_input_grad[0]:
            dx = dy @ w16
            if dx.dtype != ctx.x_dtype:
                dx = dx.to(ctx.x_dtype)
        if ctx.needs_input_grad[1]:
            # fp32 weight gradient straight out of the GEMM (aten::mm.dtype): no bf16 rounding of dW, no cast kernel
            dw = torch.mm(dy.t(), x16, out_dtype=torch.float32)
            if ctx.extra > 0:
                dw = dw[:, :dw.shape[1] - ctx.extra]
        if ctx.needs_input_grad[2] and db is None:
            db = (_ones(dy.shape[0], dy.device) @ dy)[0].float()  # column sums on the tensor cores
        return dx, dw, db, None


def mlp_forward(layers: nn.Sequential, x: torch.Tensor) -> torch.Tensor:
    """Apply ``layers`` (Linear/ReLU/Sigmoid modules of the reference MLP) through the fused bf16 path."""
    mods = list(layers)
    i = 0
    while i < len(mods):
        m = mods[i]
        if isinstance(m, (nn.Linear, PaddedLinear)):
            nxt = mods[i + 1] if i + 1 < len(mods) else None
            relu = isinstance(nxt, nn.ReLU)
            x = _LinearReLU.apply(x, m.weight, m.bias, relu)
            i += 2 if relu else 1
        else:
            x = m(x)
            i += 1
    return x
