"""Per-table sum-pooled EmbeddingBag with a sparse gradient.

Host-side mirror of the per-table seam of the reference: ``nn.EmbeddingBag(n, m, mode="sum", sparse=True)``
built in ``DLRM_Net.create_emb`` (``P/dlrm_s_pytorch.py:213-238``) and called as ``E(indices, offsets)`` in
``apply_emb`` (``P/dlrm_s_pytorch.py:299-305``); the gradient of ``E.weight`` reaches ``optimizer.step()``
(``P/dlrm_s_pytorch.py:805,1059``) as a sparse COO tensor.

Forward = ``b200_emb_fwd`` with one table; backward = ``b200_emb_bwd_coalesce`` (stable radix sort, dedup,
segment reduction) -> a COALESCED ``torch.sparse_coo_tensor`` -- the reference's ATen path hands the
optimizer the un-coalesced ``[nnz, D]`` gradient instead and leaves the de-duplication to ``add_``.
The merged module (``cloudtik_b200.merged_embedding``) is the fast path; this class exists so that code
written against ``emb_l[k](indices, offsets)`` + ``torch.optim`` keeps working unchanged on the GPU.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from . import ops


class _BagFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, weight, module, indices, offsets):
        B = offsets.numel()
        nnz = indices.numel()
        off = torch.cat([offsets, offsets.new_tensor([nnz])])  # ATen convention: no trailing offset
        out = torch.empty(B, weight.shape[1], dtype=weight.dtype, device=weight.device)
        ops.emb_fwd(module._tables(), indices, off, B, out, B * weight.shape[1], weight.shape[1])
        ctx.module = module
        ctx.save_for_backward(indices, off)
        return out

    @staticmethod
    def backward(ctx, grad):
        indices, off = ctx.saved_tensors
        m = ctx.module
        B = off.numel() - 1
        D = m.embedding_dim
        grad = grad.contiguous()
        if indices.numel() == 0:
            g = torch.sparse_coo_tensor(torch.empty(1, 0, dtype=torch.int64, device=grad.device),
                                        torch.empty(0, D, dtype=m.weight.dtype, device=grad.device),
                                        size=m.weight.shape)
            return g, None, None, None
        ws = m._workspace(indices.numel())
        rows, vals = ops.emb_bwd_coalesce(m._tables(), indices, off, B, grad, B * D, D, ws)
        g = torch.sparse_coo_tensor(rows.unsqueeze(0), vals.to(m.weight.dtype), size=m.weight.shape,
                                    is_coalesced=True)
        return g, None, None, None


class EmbeddingBag(nn.Module):
    """``EmbeddingBag(num_embeddings, embedding_dim, mode="sum", sparse=True)`` on the B200 kernels."""

    def __init__(self, num_embeddings, embedding_dim, mode="sum", sparse=True, _weight=None, device="cuda",
                 dtype=torch.float32):
        super().__init__()
        if mode != "sum" or not sparse:
            raise NotImplementedError("only mode='sum', sparse=True (the reference's configuration) is supported")
        self.num_embeddings, self.embedding_dim = int(num_embeddings), int(embedding_dim)
        if _weight is None:
            w = torch.empty(self.num_embeddings, self.embedding_dim, dtype=dtype, device=device)
            w.uniform_(-math.sqrt(1 / self.num_embeddings), math.sqrt(1 / self.num_embeddings))
        else:
            w = _weight.detach().to(device=device, dtype=dtype).contiguous()
        self.weight = nn.Parameter(w)
        self._ts = None
        self._ts_ptr = None
        self._ws = None

    def _tables(self):
        if self._ts is None or self._ts_ptr != self.weight.data_ptr():
            self._ts = ops.TableSet([self.weight.data])
            self._ts_ptr = self.weight.data_ptr()
        return self._ts

    def _workspace(self, nnz):
        need = ops.emb_bwd_workspace_bytes(nnz, self.embedding_dim)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.weight.device)
        return self._ws

    def forward(self, indices, offsets):
        if not indices.is_cuda:
            raise RuntimeError("cloudtik_b200.EmbeddingBag needs CUDA tensors (no CPU fallback)")
        return _BagFn.apply(self.weight, self, indices, offsets)
