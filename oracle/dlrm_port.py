"""torch-CPU restatement of the reference ``DLRM_Net`` training step.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  This is the model-level oracle
and the timed CPU baseline (``bench.py --impl reference`` / ``cpu_baseline``): it
calls the same ATen operators the reference calls (``nn.EmbeddingBag(sparse=True)``,
``torch.bmm``, ``nn.Linear``, ``BCELoss``, ``torch.optim.SGD``) in the same order, so
on the same seed it reproduces the reference's losses bit for bit (pinned by
``tests/test_oracle_golden.py`` against fixtures written by ``gen_golden.py``).

``P/`` = reference ``applications/ai/quickstart/models/recommendation/pytorch/dlrm/product``.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn


def build_mlp(ln, sigmoid_layer, rng):
    """``DLRM_Net.create_mlp`` ``P/dlrm_s_pytorch.py:172-211``: Linear layers with
    normal(0, sqrt(2/(m+n))) weights and normal(0, sqrt(1/m)) biases drawn from the
    numpy global stream (weights first, then bias, layer by layer); ReLU after every
    layer except ``sigmoid_layer`` which gets a Sigmoid."""
    layers = []
    for i in range(len(ln) - 1):
        n, m = int(ln[i]), int(ln[i + 1])
        lin = nn.Linear(n, m, bias=True)
        W = rng.normal(0.0, np.sqrt(2 / (m + n)), size=(m, n)).astype(np.float32)
        b = rng.normal(0.0, np.sqrt(1 / m), size=m).astype(np.float32)
        lin.weight.data = torch.tensor(W, requires_grad=True)
        lin.bias.data = torch.tensor(b, requires_grad=True)
        layers.append(lin)
        layers.append(nn.Sigmoid() if i == sigmoid_layer else nn.ReLU())
    return nn.Sequential(*layers)


def build_tables(D, ln_emb, rng, sparse=True):
    """``DLRM_Net.create_emb`` ``P/dlrm_s_pytorch.py:213-238`` with
    ``np_init_emb_weight=True``: U(-sqrt(1/n), sqrt(1/n)) rows from the numpy stream."""
    tables = nn.ModuleList()
    for n in ln_emb:
        n = int(n)
        W = rng.uniform(low=-np.sqrt(1 / n), high=np.sqrt(1 / n), size=(n, D)).astype(np.float32)
        tables.append(nn.EmbeddingBag(n, D, mode="sum", sparse=sparse,
                                      _weight=torch.from_numpy(W).requires_grad_()))
    return tables


class DLRMPort(nn.Module):
    """Single-process ``DLRM_Net`` (``P/dlrm_s_pytorch.py:171-397``), dot interaction.

    Construction order of random draws matches ``DLRM_Net.__init__`` (:240-271):
    ``l_emb_seeds`` randint, bottom MLP, top MLP, then the tables.
    ``rng`` is ``numpy.random`` (the global legacy stream, as in the reference) or a
    ``RandomState``.
    """

    def __init__(self, D, ln_emb, ln_bot, ln_top, sigmoid_bot=-1, sigmoid_top=None,
                 loss_threshold=0.0, rng=np.random):
        super().__init__()
        ln_emb = np.asarray(ln_emb)
        ln_top = np.asarray(ln_top)
        if sigmoid_top is None:
            sigmoid_top = ln_top.size - 2  # P/dlrm_s_pytorch.py:795
        self.loss_threshold = loss_threshold
        self.l_emb_seeds = rng.randint(low=0, high=100000, size=len(ln_emb))
        self.bot_l = build_mlp(np.asarray(ln_bot), sigmoid_bot, rng)
        self.top_l = build_mlp(ln_top, sigmoid_top, rng)
        self.emb_l = build_tables(int(D), ln_emb, rng)
        self.loss_fn = nn.BCELoss(reduction="mean")

    def apply_emb(self, lS_o, lS_i):
        # P/dlrm_s_pytorch.py:290-307
        return [E(lS_i[k], lS_o[k]) for k, E in enumerate(self.emb_l)]

    @staticmethod
    def interact_features(x, ly):
        # P/dlrm_s_pytorch.py:314-333
        B, d = x.shape
        T = torch.cat([x] + list(ly), dim=1).view((B, -1, d))
        Z = torch.bmm(T, torch.transpose(T, 1, 2))
        _, ni, nj = Z.shape
        li = torch.tensor([i for i in range(ni) for j in range(i)])
        lj = torch.tensor([j for i in range(nj) for j in range(i)])
        return torch.cat([x, Z[:, li, lj]], dim=1)

    def forward(self, dense_x, lS_o, lS_i):
        # sequential_forward, P/dlrm_s_pytorch.py:372-397
        x = self.bot_l(dense_x)
        ly = self.apply_emb(lS_o, lS_i)
        z = self.interact_features(x, ly)
        p = self.top_l(z)
        if 0.0 < self.loss_threshold < 1.0:
            p = torch.clamp(p, min=self.loss_threshold, max=1.0 - self.loss_threshold)
        return p


def train_step(model, optimizer, X, lS_o, lS_i, T, bf16=False):
    """One iteration of the reference loop ``P/dlrm_s_pytorch.py:1029-1059``:
    forward (optionally under CPU autocast), BCE loss, ``zero_grad(set_to_none)``,
    backward, ``optimizer.step()``.  Returns the loss as a python float (the
    reference's per-step ``E.detach().cpu().numpy()`` at :1048)."""
    with torch.autocast("cpu", dtype=torch.bfloat16, enabled=bf16):
        Z = model(X, lS_o, lS_i).float()
    E = model.loss_fn(Z, T)
    L = float(E.detach().cpu().numpy())
    optimizer.zero_grad(set_to_none=True)
    E.backward()
    optimizer.step()
    return L


def uniform_batch(rs, B, ln_emb, pooling=1, n_dense=13, unique_in_bag=False):
    """Synthetic Criteo-shaped batch: dense ``U[0,1)``, targets
    ``round(U[0,1))``, indices ``floor(U*n)`` uniform over rows, fixed pooling factor.
    With ``unique_in_bag`` the per-bag ``np.unique`` of the reference generator
    (``T/dlrm_data_pytorch.py:791-799``) is applied, which makes bags ragged."""
    X = rs.rand(B, n_dense).astype(np.float32)
    T = np.round(rs.rand(B, 1)).astype(np.float32)
    lS_o, lS_i = [], []
    for n in ln_emb:
        n = int(n)
        idx = np.floor(rs.rand(B, pooling) * n).astype(np.int64)
        if unique_in_bag:
            bags = [np.unique(r) for r in idx]
            lens = np.array([len(b) for b in bags], dtype=np.int64)
            lS_i.append(np.concatenate(bags))
            lS_o.append(np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64))
        else:
            lS_i.append(idx.reshape(-1))
            lS_o.append(np.arange(B, dtype=np.int64) * pooling)
    return X, lS_o, lS_i, T
