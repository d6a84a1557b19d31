"""Pairwise dot-product feature interaction (differentiable).

Host-side mirror of ``DLRM_Net.interact_features`` (reference ``P/dlrm_s_pytorch.py:309-334``,
``T/dlrm_s_pytorch.py:407-444``) with the call shape of the reference's own fused seam
``ipex.nn.functional.interaction(*T)`` (``P/dlrm_s_pytorch.py:310-312``): ``interaction(x, *ly)``
returns ``[B, D + F(F-1)/2]`` = ``[x, <T_i, T_j> for i in 1..F-1 for j in 0..i-1]`` with
``T = [x] + ly``.

When ``ly`` are the views returned by ``MergedEmbeddingBagWithSGD`` the pooled rows already sit
in a ``[B, F, D]`` buffer with slot 0 free: x is copied into slot 0 and the kernel reads the
buffer in place (no ``torch.cat`` pass).  Backward writes ``dT [B, F, D]`` once and hands out
views, which the merged embedding's backward consumes in place as well.
"""
from __future__ import annotations

import torch

from . import ops
from .merged_embedding import lookup_pooled


def _pooled_base(x: torch.Tensor, ly):
    base = lookup_pooled(ly[0])
    if base is None:
        return None
    B, F, D = base.shape
    if F != len(ly) + 1 or tuple(x.shape) != (B, D) or not base.is_contiguous():
        return None
    st = base.untyped_storage().data_ptr()
    for k, t in enumerate(ly):
        if (t.untyped_storage().data_ptr() != st or t.dtype != base.dtype or tuple(t.shape) != (B, D)
                or t.stride() != (F * D, 1) or t.storage_offset() != base.storage_offset() + (1 + k) * D):
            return None
    return base


class _InteractionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pad_to, x, *ly):
        itself = pad_to < 0  # sign bit carries the arch_interaction_itself option
        pad_to = abs(pad_to)
        ctx.itself = itself
        base = _pooled_base(x, ly)
        if base is not None:
            T = base
            T[:, 0].copy_(x)
        else:
            dt = ly[0].dtype if ly else x.dtype
            T = torch.stack([x.to(dt)] + [t.to(dt) for t in ly], dim=1).contiguous()
        ctx.save_for_backward(T)
        ctx.x_dtype = x.dtype
        ctx.ly_dtypes = [t.dtype for t in ly]
        return ops.interact_fwd(T, pad_to=pad_to, itself=itself)

    @staticmethod
    def backward(ctx, dR):
        (T,) = ctx.saved_tensors
        dT = ops.interact_bwd(T, dR.contiguous().to(T.dtype), itself=ctx.itself)
        dx = dT[:, 0]
        if dx.dtype != ctx.x_dtype:
            dx = dx.to(ctx.x_dtype)
        dly = []
        for k, dt in enumerate(ctx.ly_dtypes):
            g = dT[:, 1 + k]
            dly.append(g if g.dtype == dt else g.to(dt))
        return (None, dx, *dly)


class _InteractionShardedFn(torch.autograd.Function):
    """Distributed variant: ``chunks[r]`` is what the pooled-embedding all-to-all delivered from rank r,
    a contiguous ``[lN, lS_r * D]`` block.  Forward packs x and the chunks into ``T [lN, F, D]`` with one
    strided copy per source rank (instead of F slices + stack), backward unpacks ``dT`` the same way, so
    autograd never materialises per-feature slices."""

    @staticmethod
    def forward(ctx, pad_to, x, *chunks):
        lN, D = x.shape
        counts = [c.shape[1] // D for c in chunks]
        F = 1 + sum(counts)
        dt = chunks[0].dtype
        T = torch.empty(lN, F, D, dtype=dt, device=x.device)
        T[:, 0].copy_(x)
        f0 = 1
        for c, n in zip(chunks, counts):
            T[:, f0:f0 + n].copy_(c.view(lN, n, D))
            f0 += n
        ctx.save_for_backward(T)
        ctx.counts = counts
        ctx.x_dtype = x.dtype
        return ops.interact_fwd(T, pad_to=pad_to)

    @staticmethod
    def backward(ctx, dR):
        (T,) = ctx.saved_tensors
        lN, F, D = T.shape
        dT = ops.interact_bwd(T, dR.contiguous().to(T.dtype))
        dx = dT[:, 0].to(ctx.x_dtype)
        grads, f0 = [], 1
        for n in ctx.counts:
            grads.append(dT[:, f0:f0 + n].reshape(lN, n * D))  # one contiguous copy per destination rank
            f0 += n
        return (None, dx, *grads)


def interaction_sharded(x: torch.Tensor, chunks, pad_to: int = 1) -> torch.Tensor:
    """Interaction of the local-batch dense feature ``x [lN, D]`` with the per-source-rank pooled
    chunks ``[lN, lS_r * D]`` returned by ``ext_dist.alltoall(...).wait()`` (features in rank order =
    global table order for the contiguous table partition)."""
    return _InteractionShardedFn.apply(int(pad_to), x, *chunks)


def interaction_cat(x: torch.Tensor, *ly: torch.Tensor) -> torch.Tensor:
    """``--arch-interaction-op=cat`` (``T/dlrm_s_pytorch.py:434-436``): ``R = cat([x] + ly, dim=1)``.  When ``ly`` are
    the merged embedding's views, x is copied into the free slot of the pooled ``[B, F, D]`` buffer and the result
    is that buffer seen as ``[B, F*D]`` -- no concatenation pass; otherwise a plain ``torch.cat``."""
    base = _pooled_base(x, ly)
    if base is None or torch.is_grad_enabled() and (x.requires_grad or any(t.requires_grad for t in ly)):
        return torch.cat([x.to(ly[0].dtype)] + list(ly), dim=1)  # autograd routes the gradient slices
    base[:, 0].copy_(x)
    return base.view(base.shape[0], -1)


def interaction(x: torch.Tensor, *ly: torch.Tensor, pad_to: int = 1, itself: bool = False) -> torch.Tensor:
    """``interaction(x, *ly)`` -- see module docstring.  CUDA tensors only.

    ``itself`` (``--arch-interaction-itself``, ``T/dlrm_s_pytorch.py:428``): pairs (i, j <= i), the output is
    ``[B, D + F(F+1)/2]``.

    ``pad_to`` > 1 rounds the output width up to a multiple of ``pad_to`` with zero columns (e.g.
    479 -> 480) so that the GEMM consuming it gets an aligned K; ``PaddedLinear`` accepts that."""
    if not x.is_cuda:
        raise RuntimeError("cloudtik_b200.interaction needs CUDA tensors (no CPU fallback)")
    return _InteractionFn.apply(-int(pad_to) if itself else int(pad_to), x, *ly)


class PaddedLinear(torch.nn.Linear):
    """``nn.Linear`` (same parameters, same state_dict keys) that also accepts an input with extra
    zero columns on the right: the weight is zero-padded on the fly.  Used for the first top-MLP
    layer (reference P/dlrm_s_pytorch.py:172-211 builds a plain Linear(479, 1024)): K = 479 sends
    cuBLAS to an unaligned legacy kernel ~6x slower than K = 480."""

    def forward(self, input):
        extra = input.shape[-1] - self.in_features
        if extra == 0:
            return torch.nn.functional.linear(input, self.weight, self.bias)
        if extra < 0:
            raise RuntimeError("PaddedLinear: input narrower than in_features")
        return torch.nn.functional.linear(input, torch.nn.functional.pad(self.weight, (0, extra)), self.bias)
