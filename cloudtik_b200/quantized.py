"""int8 embedding tables for inference.

The reference's ``--int8`` inference path (``P/dlrm_s_pytorch.py:428-447,868-883``) prepares the model with
``QConfig(activation=MinMaxObserver(per_tensor_symmetric, qint8), weight=PerChannelMinMaxObserver(qint8,
per_channel_symmetric))`` and converts it through IPEX; for an ``EmbeddingBag`` that means a symmetric int8
weight with zero point 0.  Here the table part of that is done natively: ``QuantizedMergedEmbeddingBag`` holds
every table as int8 ``[rows, D]`` + one fp32 scale per row (the observer's channel axis 0) and its forward is
``b200_emb_fwd_q8`` -- a quarter of the fp32 gather bytes, dequantised in registers, written in the same
``[B, T+1, D]`` layout the interaction kernel reads in place.  The MLPs stay in bf16/fp32 on torch.
"""
from __future__ import annotations

from typing import List

import torch
import torch.nn as nn

from . import ops
from .merged_embedding import _register_pooled

_EPS = float(torch.finfo(torch.float32).eps)


def quantize_rowwise_int8(w: torch.Tensor, chunk_rows: int = 2_000_000):
    """(q int8 [n, D], scale fp32 [n]): PerChannelMinMaxObserver(ch_axis=0, qint8, per_channel_symmetric) +
    quantize_per_channel, in fp32 on the device (bit-exact with ``oracle.ops_numpy.quantize_rowwise_int8``).
    Done in row chunks so that a 20 GB table needs ~3 GB of temporaries, not three copies of itself."""
    n = w.shape[0]
    q = torch.empty(w.shape, dtype=torch.int8, device=w.device)
    scale = torch.empty(n, dtype=torch.float32, device=w.device)
    for a in range(0, n, chunk_rows):
        c = w[a:a + chunk_rows].detach().float()
        mn = torch.clamp(c.amin(dim=1), max=0.0)
        mx = torch.clamp(c.amax(dim=1), min=0.0)
        # tensor / tensor: a scalar divisor would be turned into a multiplication by its reciprocal on CUDA
        sc = torch.clamp(torch.maximum(-mn, mx) / torch.full_like(mn, 127.5), min=_EPS)
        inv = torch.ones_like(sc) / sc
        q[a:a + chunk_rows] = torch.clamp(torch.round(c * inv[:, None]), -128, 127).to(torch.int8)
        scale[a:a + chunk_rows] = sc
    return q, scale


class QuantizedMergedEmbeddingBag(nn.Module):
    """Forward-only merged EmbeddingBag over int8 tables; same call convention as
    ``MergedEmbeddingBagWithSGD`` (``(indices, offsets, _)``, need_linearize flag, ``.weights``, ``.n_tables``)."""

    def __init__(self, qweights: List[torch.Tensor], scales: List[torch.Tensor], output_dtype=torch.float32):
        super().__init__()
        self.n_tables = len(qweights)
        self.dim = int(qweights[0].shape[1])
        for k, (q, s) in enumerate(zip(qweights, scales)):
            self.register_buffer(f"qweight_{k}", q)
            self.register_buffer(f"scale_{k}", s)
        self.output_dtype = output_dtype
        self.reserve_dense_slot = True
        self.error_flag = None
        self.profile = None
        self.tables = ops.QuantTableSet(self.qweights, self.scales)
        rows = self.tables.rows
        base = [0]
        for r in rows:
            base.append(base[-1] + r)
        self._row_base = base

    @classmethod
    def from_float(cls, emb, output_dtype=None):
        """Quantise a ``MergedEmbeddingBagWithSGD`` (fp32 tables; bf16 tables go through their fp32 masters)."""
        qs = [quantize_rowwise_int8(w) for w in emb.master_weights()]
        m = cls([q for q, _ in qs], [s for _, s in qs], output_dtype or emb.output_dtype or torch.float32)
        m.reserve_dense_slot = emb.reserve_dense_slot
        return m

    @property
    def qweights(self):
        return [getattr(self, f"qweight_{k}") for k in range(self.n_tables)]

    @property
    def scales(self):
        return [getattr(self, f"scale_{k}") for k in range(self.n_tables)]

    @property
    def weights(self):
        """Dequantised fp32 tables (for inspection; the kernel never materialises them)."""
        return [q.float() * s[:, None] for q, s in zip(self.qweights, self.scales)]

    def _apply(self, fn, *a, **k):
        r = super()._apply(fn, *a, **k)
        if getattr(self, "tables", None) is not None:
            self.tables = ops.QuantTableSet(self.qweights, self.scales)
        return r

    def linearize_indices_and_offsets(self, indices, offsets, include_last_offsets):
        dev = self.qweights[0].device
        idx = [i.to(dev, torch.int64) for i in indices]
        offs = []
        for o, inc in zip(offsets, include_last_offsets):
            o = o.to(dev, torch.int64)
            offs.append(o[:-1] if inc else o)
        nnz = [0]
        for i in idx:
            nnz.append(nnz[-1] + i.numel())
        lin_idx = torch.cat(idx)
        lin_off = torch.cat([o + nnz[t] for t, o in enumerate(offs)]
                            + [torch.tensor([nnz[-1]], dtype=torch.int64, device=dev)])
        iwro = torch.cat([i + self._row_base[t] for t, i in enumerate(idx)])
        return lin_idx, lin_off, iwro

    def forward(self, input, need_linearize_indices_and_offsets=None):
        need = True if need_linearize_indices_and_offsets is None else bool(need_linearize_indices_and_offsets)
        if need:
            inc = input[2] if len(input) > 2 else [False] * len(input[0])
            indices, offsets, _ = self.linearize_indices_and_offsets(input[0], input[1], inc)
        else:
            indices, offsets = input[0], input[1]
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            raise RuntimeError("int8 tables are inference-only")
        T, D = self.n_tables, self.dim
        B = (offsets.numel() - 1) // T
        s0 = 1 if self.reserve_dense_slot else 0
        buf = torch.empty(B, T + s0, D, dtype=self.output_dtype, device=indices.device)
        if s0:
            _register_pooled(buf)
        ops.emb_fwd_q8(self.tables, indices, offsets, B, buf[:, s0:], D, (T + s0) * D, self.error_flag)
        return [buf[:, s0 + k] for k in range(T)]


def quantize_embeddings_int8(dlrm):
    """Swap ``dlrm.emb_l`` for its int8 counterpart (in place) and put the model in eval mode: the table half of
    the reference's ``--int8`` convert step.  Returns the model."""
    dlrm.emb_l = QuantizedMergedEmbeddingBag.from_float(dlrm.emb_l)
    dlrm.eval()
    return dlrm
