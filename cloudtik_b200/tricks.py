"""Compositional / mixed-dimension embedding operators on the B200 EmbeddingBag kernels.

Mirrors of the reference's ``tricks`` package (``T/tricks/qr_embedding_bag.py``, ``T/tricks/md_embedding_bag.py``),
which ``DLRM_Net.create_emb`` instantiates behind ``--qr-flag`` / ``--md-flag`` (``T/dlrm_s_pytorch.py:246-252``):

* ``QREmbeddingBag`` (``qr_embedding_bag.py:25-178``): quotient-remainder trick -- two small tables, indexed with
  ``idx / num_collisions`` and ``idx % num_collisions``, each sum-pooled, the two POOLED results combined by
  ``concat`` / ``add`` / ``mult``.  Each lookup is ``cloudtik_b200.embedding_bag.EmbeddingBag`` (``b200_emb_fwd``
  forward, ``b200_emb_bwd_coalesce`` sparse gradient); the combine and its gradient are plain elementwise torch.
  The reference computes the quotient as ``(input / num_collisions).long()`` -- a TRUE division, i.e. through
  float32 -- so ids above 2**24 can land on a neighbouring row; the mirror keeps that arithmetic (quirk, not fixed).
* ``PrEmbeddingBag`` (``md_embedding_bag.py:63-81``): a narrow table followed by a bias-free projection to the
  base dimension; ``md_solver`` / ``alpha_power_rule`` / ``pow_2_round`` (``:20-60``) pick the dimensions.

Only ``mode="sum"``, ``sparse=True`` -- the configuration the reference's DLRM uses.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from .embedding_bag import EmbeddingBag


class QREmbeddingBag(nn.Module):
    def __init__(self, num_categories, embedding_dim, num_collisions, operation="mult", mode="sum", sparse=True,
                 _weight=None, device="cuda"):
        super().__init__()
        assert operation in ["concat", "mult", "add"], "Not valid operation!"
        if mode != "sum" or not sparse:
            raise NotImplementedError("only mode='sum', sparse=True (what DLRM_Net.create_emb passes) is supported")
        self.num_categories = num_categories
        if isinstance(embedding_dim, int) or len(embedding_dim) == 1:
            self.embedding_dim = [embedding_dim, embedding_dim]
        else:
            self.embedding_dim = list(embedding_dim)
        self.num_collisions = num_collisions
        self.operation = operation
        if operation in ("add", "mult"):
            assert self.embedding_dim[0] == self.embedding_dim[1], "Embedding dimensions do not match!"
        self.num_embeddings = [int(np.ceil(num_categories / num_collisions)), num_collisions]
        if _weight is None:
            wq = torch.empty(self.num_embeddings[0], self.embedding_dim[0]).uniform_(np.sqrt(1 / num_categories), 1.0)
            wr = torch.empty(self.num_embeddings[1], self.embedding_dim[1]).uniform_(np.sqrt(1 / num_categories), 1.0)
        else:
            assert list(_weight[0].shape) == [self.num_embeddings[0], self.embedding_dim[0]]
            assert list(_weight[1].shape) == [self.num_embeddings[1], self.embedding_dim[1]]
            wq, wr = _weight
        self.emb_q = EmbeddingBag(self.num_embeddings[0], self.embedding_dim[0], _weight=wq, device=device)
        self.emb_r = EmbeddingBag(self.num_embeddings[1], self.embedding_dim[1], _weight=wr, device=device)
        self.mode, self.sparse = mode, sparse

    @property
    def weight_q(self):
        return self.emb_q.weight

    @property
    def weight_r(self):
        return self.emb_r.weight

    def forward(self, input, offsets=None, per_sample_weights=None):
        if per_sample_weights is not None:
            raise NotImplementedError("per_sample_weights is not on the reference's DLRM path")
        # true division through float32, like the reference on the CPU.  Tensor / tensor on purpose: with a Python
        # scalar divisor CUDA multiplies by the reciprocal, which rounds differently from the CPU's IEEE division
        f = input.to(torch.float32)
        input_q = (f / torch.full_like(f, float(self.num_collisions))).long()
        input_r = torch.remainder(input, self.num_collisions).long()
        embed_q = self.emb_q(input_q, offsets)
        embed_r = self.emb_r(input_r, offsets)
        if self.operation == "concat":
            return torch.cat((embed_q, embed_r), dim=1)
        if self.operation == "add":
            return embed_q + embed_r
        return embed_q * embed_r


def md_solver(n, alpha, d0=None, B=None, round_dim=True, k=None):
    """``md_embedding_bag.py:20-41``: embedding dimension per table from its row count (popularity-weighted when
    ``k`` gives the queries per inference), by the alpha power rule; dims come back in the caller's table order."""
    n_sorted, perm = torch.sort(n)
    weight = torch.ones(n_sorted.numel()) if k is None else k[perm]
    dims = alpha_power_rule(n_sorted.to(torch.float) / weight, alpha, d0=d0, B=B)
    if round_dim:
        dims = pow_2_round(dims)
    back = torch.empty_like(perm)
    back[perm] = torch.arange(perm.numel())
    return dims[back]


def alpha_power_rule(n, alpha, d0=None, B=None):
    """``md_embedding_bag.py:43-56``: d_i = lambda * n_i^-alpha, at least 1, rounded; lambda from the base
    dimension ``d0`` of the smallest table (which then keeps exactly ``d0``) or from a parameter budget ``B``."""
    nf = n.to(torch.float)
    if d0 is not None:
        lam = d0 * nf[0] ** alpha
    elif B is not None:
        lam = B / (nf ** (1 - alpha)).sum()
    else:
        raise ValueError("Must specify either d0 or B")
    dims = torch.clamp(torch.ones(nf.numel()) * lam * nf ** (-alpha), min=1.0)
    if d0 is not None:
        dims[0] = d0
    return torch.round(dims).to(torch.long)


def pow_2_round(dims):
    """``md_embedding_bag.py:59-60``: nearest power of two (in log space)."""
    return torch.pow(2.0, torch.round(torch.log2(dims.to(torch.float))))


class PrEmbeddingBag(nn.Module):
    """Narrow table + projection to ``base_dim`` (``md_embedding_bag.py:63-81``)."""

    def __init__(self, num_embeddings, embedding_dim, base_dim, device="cuda"):
        super().__init__()
        w = torch.empty(num_embeddings, embedding_dim)
        torch.nn.init.xavier_uniform_(w)
        self.embs = EmbeddingBag(num_embeddings, embedding_dim, _weight=w, device=device)
        if embedding_dim < base_dim:
            self.proj = nn.Linear(embedding_dim, base_dim, bias=False, device=device)
            torch.nn.init.xavier_uniform_(self.proj.weight)
        elif embedding_dim == base_dim:
            self.proj = nn.Identity()
        else:
            raise ValueError("Embedding dim " + str(embedding_dim) + " > base dim " + str(base_dim))

    def forward(self, input, offsets=None, per_sample_weights=None):
        if per_sample_weights is not None:
            raise NotImplementedError("per_sample_weights is not on the reference's DLRM path")
        return self.proj(self.embs(input, offsets))
