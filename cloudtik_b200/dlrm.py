"""DLRM model wired to the B200 kernels.

Mirror of the reference ``DLRM_Net`` (``P/dlrm_s_pytorch.py:171-397``): same constructor
arguments, same weight initialisation order from the numpy stream (so a seed gives the same
model as the reference), same ``forward(dense_x, *emb_args)`` dispatch.  The dense MLPs stay
``torch.nn`` (cuBLAS); the embedding tables are a ``MergedEmbeddingBagWithSGD`` (pool + fused
sparse update) and the interaction is ``cloudtik_b200.interaction`` -- the two seams the
reference itself switches on with ``--ipex-merged-emb`` / ``--ipex-interaction``.
"""
from __future__ import annotations

import sys

import numpy as np
import torch
import torch.nn as nn

from . import ext_dist
from .interaction import PaddedLinear, interaction, interaction_sharded
from .merged_embedding import MergedEmbeddingBagWithSGD
from .mlp import mlp_forward


class DLRM_Net(nn.Module):
    def create_mlp(self, ln, sigmoid_layer, pad_first=False):
        # P/dlrm_s_pytorch.py:172-211 (pad_first: first layer tolerates zero-padded input columns)
        layers = nn.ModuleList()
        for i in range(0, ln.size - 1):
            n, m = int(ln[i]), int(ln[i + 1])
            LL = PaddedLinear(n, m, bias=True) if (pad_first and i == 0) else nn.Linear(n, m, bias=True)
            W = np.random.normal(0.0, np.sqrt(2 / (m + n)), size=(m, n)).astype(np.float32)
            bt = np.random.normal(0.0, np.sqrt(1 / m), size=m).astype(np.float32)
            LL.weight.data = torch.tensor(W, requires_grad=True)
            LL.bias.data = torch.tensor(bt, requires_grad=True)
            layers.append(LL)
            layers.append(nn.Sigmoid() if i == sigmoid_layer else nn.ReLU())
        return torch.nn.Sequential(*layers)

    def create_emb(self, m, ln, local_ln_emb=None, np_init_emb_weight=False, device="cuda", lr=0.01,
                   optimizer="sgd"):
        # P/dlrm_s_pytorch.py:213-238; np_init_emb_weight=True draws every table from the numpy
        # stream on the host (parity runs); otherwise rows are drawn on the device with the same
        # U(-sqrt(1/n), sqrt(1/n)) law (Terabyte-size tables never exist on the host).
        weights = []
        ids = range(ln.size) if local_ln_emb is None else local_ln_emb
        for i in ids:
            n = int(ln[i])
            if np_init_emb_weight:
                W = np.random.uniform(low=-np.sqrt(1 / n), high=np.sqrt(1 / n), size=(n, m)).astype(np.float32)
                weights.append(torch.from_numpy(W).to(device))
            else:
                w = torch.empty(n, m, dtype=torch.float32, device=device)
                w.uniform_(-float(np.sqrt(1 / n)), float(np.sqrt(1 / n)))
                weights.append(w)
        return MergedEmbeddingBagWithSGD(weights, lr=lr, optimizer=optimizer)

    def __init__(self, m_spa=None, ln_emb=None, ln_bot=None, ln_top=None, sigmoid_bot=-1, sigmoid_top=-1,
                 weighted_pooling=None, loss_threshold=0.0, np_init_emb_weight=False, device="cuda",
                 lr=0.01, emb_optimizer="sgd", bf16=False, n_rowwise=0, distributed=None):
        """``distributed=False`` builds the single-process model (all tables local) even inside an initialised
        process group -- the oracle of the multi-GPU parity checks runs that way next to the sharded model."""
        super().__init__()
        self.world = ext_dist.my_size if (distributed is None or distributed) and ext_dist.my_size > 1 else 1
        if weighted_pooling is not None:
            raise NotImplementedError("weighted pooling is not on the reference's product path")
        self.loss_threshold = loss_threshold
        self.bf16 = bf16
        self.fused_mlp = True
        self.peer_exchange = None  # set by enable_peer_exchange(): fused a2a over NVLink peer memory
        self.peer_exchanges = None  # one per micro-batch of a step (peer_exchange is the first)
        ln_emb = np.asarray(ln_emb)
        # the LAST n_rowwise tables are row-wise sharded over the ranks (hot tables, config 5); the others
        # are partitioned table-wise exactly like the reference (P/dlrm_s_pytorch.py:256-264)
        self.n_rowwise = int(n_rowwise) if self.world > 1 else 0
        self.rowwise = None
        if self.world > 1:
            n_emb = len(ln_emb) - self.n_rowwise
            self.n_global_emb = len(ln_emb)
            self.rank = ext_dist.dist.get_rank()
            self.ln_emb = [i for i in range(n_emb)]
            self.n_local_emb, self.n_emb_per_rank = ext_dist.get_split_lengths(n_emb)
            self.local_ln_emb_slice = ext_dist.get_my_slice(n_emb)
            self.local_ln_emb = self.ln_emb[self.local_ln_emb_slice]
        else:
            self.local_ln_emb = None
        self.l_emb_seeds = np.random.randint(low=0, high=100000, size=len(ln_emb))
        self.bot_l = self.create_mlp(np.asarray(ln_bot), sigmoid_bot, pad_first=True).to(device)
        self.dense_pad = (-int(ln_bot[0])) % 8  # 13 dense features -> 16 zero-padded columns (aligned GEMM K)
        self.top_l = self.create_mlp(np.asarray(ln_top), sigmoid_top, pad_first=True).to(device)
        self.interaction_pad = 8  # interaction output rounded up to a multiple of 8 columns
        self.emb_l = self.create_emb(m_spa, ln_emb, self.local_ln_emb, np_init_emb_weight, device, lr,
                                     emb_optimizer)
        if self.n_rowwise:
            from .rowwise import RowWiseShardedTables
            hot = [int(n) for n in ln_emb[len(ln_emb) - self.n_rowwise:]]
            self.rowwise = RowWiseShardedTables(hot, m_spa, ext_dist.my_size, self.rank, device, lr=lr,
                                                optimizer=emb_optimizer,
                                                out_dtype=torch.bfloat16 if bf16 else None)
        if bf16:
            self.emb_l.output_dtype = torch.bfloat16
        if self.world > 1:
            self.emb_l.reserve_dense_slot = False  # pooled rows go straight into the a2a send layout
        self.need_linearize_indices_and_offsets = torch.BoolTensor([False])
        self.loss_fn = torch.nn.BCELoss(reduction="mean")

    def apply_mlp(self, x, layers):
        # bf16: fused Linear+ReLU path over the same modules (cloudtik_b200/mlp.py); a DDP-wrapped MLP
        # must go through its own forward (gradient hooks), fp32 stays plain torch
        if self.bf16 and self.fused_mlp and isinstance(layers, nn.Sequential) and x.is_cuda:
            return mlp_forward(layers, x)
        return layers(x)

    def _pad_dense(self, dense_x):
        return torch.nn.functional.pad(dense_x, (0, self.dense_pad)) if self.dense_pad else dense_x

    def apply_emb(self, emb_l, *emb_args):
        # merged call style (P/dlrm_s_pytorch.py:288-289): (indices, offsets, indices_with_row_offsets)
        if len(emb_args) == 3:
            return emb_l(emb_args, self.need_linearize_indices_and_offsets)
        # list style (P/dlrm_s_pytorch.py:290-305): (lS_o, lS_i) -> linearise on the fly
        lS_o, lS_i = emb_args
        return emb_l((list(lS_i), list(lS_o), [False] * len(lS_i)), torch.BoolTensor([True]))

    def interact_features(self, x, ly):
        return interaction(x, *ly, pad_to=self.interaction_pad)

    def forward(self, dense_x, *emb_args, chunk=0):
        """``chunk``: which micro-batch of a step cut by ``enable_peer_exchange(..., n_chunks)`` this call is
        (0 when the step is not cut)."""
        if self.world > 1:
            return self.distributed_forward(dense_x, *emb_args, chunk=chunk)
        return self.sequential_forward(dense_x, *emb_args)

    def begin_step(self, chunk_args):
        """Micro-batched peer-memory step: queue EVERY chunk's pooled forward + exchange (and index stage) on the
        embedding's side stream before the first chunk's bottom MLP runs.  ``chunk_args``: one (indices, offsets)
        pair of linearised GLOBAL-micro-batch bags per chunk.  No-op for other configurations."""
        if self.peer_exchanges is None or not (self.training and torch.is_grad_enabled()):
            return
        if len(chunk_args) != len(self.peer_exchanges):
            raise RuntimeError("begin_step: %d chunks given, the peer exchange was built for %d"
                               % (len(chunk_args), len(self.peer_exchanges)))
        from .p2p import prelaunch
        prelaunch(self.emb_l, self.peer_exchanges, [a[0] for a in chunk_args], [a[1] for a in chunk_args],
                  1 + self.local_ln_emb_slice.start)

    def _clamp(self, p):
        if 0.0 < self.loss_threshold < 1.0:
            return torch.clamp(p, min=self.loss_threshold, max=(1.0 - self.loss_threshold))
        return p

    def sequential_forward(self, dense_x, *emb_args):
        # P/dlrm_s_pytorch.py:372-397
        if len(emb_args) == 3:  # linearised bags: the pooled lookup starts now, next to the bottom MLP
            self.emb_l.prelaunch_forward(emb_args[0], emb_args[1])
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.bf16):
            x = self.apply_mlp(self._pad_dense(dense_x), self.bot_l)
            ly = self.apply_emb(self.emb_l, *emb_args)
            z = self.interact_features(x, ly)
            p = self.apply_mlp(z, self.top_l)
        return self._clamp(p)

    def distributed_forward(self, dense_x, *emb_args, chunk=0):
        # P/dlrm_s_pytorch.py:342-369: local tables on the GLOBAL batch -> all-to-all -> local batch
        batch_size = dense_x.size()[0]
        if batch_size < ext_dist.my_size:
            sys.exit("ERROR: batch_size (%d) must be larger than number of ranks (%d)"
                     % (batch_size, ext_dist.my_size))
        if self.peer_exchange is not None:
            return self._p2p_forward(dense_x, *emb_args, chunk=chunk)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.bf16):
            ly = self.apply_emb(self.emb_l, *emb_args)
            a2a_req = ext_dist.alltoall(ly, self.n_emb_per_rank)
            x = self.apply_mlp(self._pad_dense(dense_x), self.bot_l)
            ly = a2a_req.wait()
            # the reference re-slices every rank's block into per-table [lN, D] views
            # (P/dlrm_s_pytorch.py:355-357); here the blocks go to the interaction as they are
            z = interaction_sharded(x, ly, pad_to=self.interaction_pad)
            p = self.apply_mlp(z, self.top_l)
        return self._clamp(p)


def _p2p_forward(self, dense_x, *emb_args, chunk=0):
    """distributed_forward with the all-to-all fused into the embedding kernels (cloudtik_b200.p2p)."""
    from .p2p import prelaunch, sharded_emb_interact
    exch = self.peer_exchanges[chunk]
    rw_i = rw_o = None
    if len(emb_args) >= 3:
        indices, offsets = emb_args[0], emb_args[1]
        if len(emb_args) == 5:  # (+ linearised GLOBAL-batch bags of the row-wise sharded hot tables)
            rw_i, rw_o = emb_args[3], emb_args[4]
    else:
        lS_o, lS_i = emb_args
        indices, offsets, _ = self.emb_l.linearize_indices_and_offsets(list(lS_i), list(lS_o), [False] * len(lS_i))
    if (self.rowwise is not None) != (rw_i is not None):
        raise RuntimeError("row-wise sharded tables need (indices, offsets, None, rw_indices, rw_offsets)")
    feat0 = 1 + self.local_ln_emb_slice.start
    train = self.training and torch.is_grad_enabled()
    if train and len(self.peer_exchanges) == 1:
        prelaunch(self.emb_l, exch, indices, offsets, feat0)  # NVLink exchange || bottom MLP
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.bf16):
        x = self.apply_mlp(self._pad_dense(dense_x), self.bot_l)
        z = sharded_emb_interact(x, self.emb_l, exch, indices, offsets, feat0,
                                 pad_to=self.interaction_pad, rowwise=self.rowwise, rw_indices=rw_i,
                                 rw_offsets=rw_o, train=train)
        p = self.apply_mlp(z, self.top_l)
    return self._clamp(p)


DLRM_Net._p2p_forward = _p2p_forward


def enable_peer_exchange(dlrm, local_batch: int, n_chunks: int = 1):
    """Switch a table-wise sharded DLRM_Net (one process per GPU on one NVLink box) to the fused
    peer-memory exchange; ``local_batch`` = samples per rank (fixed: the buffers are symmetric).
    ``n_chunks`` > 1 cuts every step into that many micro-batches (``local_batch`` must divide), each with its own
    exchange buffers, so that one chunk's NVLink traffic runs under another chunk's compute (cloudtik_b200.p2p)."""
    from .p2p import PeerExchange
    if n_chunks < 1 or local_batch % n_chunks:
        raise ValueError("enable_peer_exchange: the local batch must be a multiple of n_chunks")
    if n_chunks > 1 and dlrm.rowwise is not None:
        raise NotImplementedError("micro-batched steps with row-wise sharded tables")
    n_feat = 1 + dlrm.n_global_emb
    dt = torch.bfloat16 if dlrm.bf16 else dlrm.emb_l.tables.dtype
    dlrm.peer_exchanges = [PeerExchange(local_batch // n_chunks, n_feat, dlrm.emb_l.dim, dt,
                                        dlrm.emb_l.tables.device, chunk=c) for c in range(n_chunks)]
    dlrm.peer_exchange = dlrm.peer_exchanges[0]
    return dlrm.peer_exchange


def train_step(dlrm, optimizer, X, emb_args, T, lr=None):
    """One iteration of the reference loop (P/dlrm_s_pytorch.py:1027-1059).  Returns the loss
    TENSOR (no host sync here; the reference's per-step ``.cpu()`` is the caller's choice)."""
    if lr is not None:
        dlrm.emb_l.sgd_args = dlrm.emb_l.sgd_args._replace(lr=lr)
    Z = dlrm(X, *emb_args).float()
    E = dlrm.loss_fn(Z, T)
    optimizer.zero_grad(set_to_none=True)
    E.backward()
    optimizer.step()
    return E.detach()


def train_step_chunked(dlrm, optimizer, batch, post_backward=None):
    """One iteration on a batch given as the flat tuple ``(X, Y, indices_0, offsets_0[, indices_1, offsets_1, ...])``:
    one (indices, offsets) pair of linearised bags per micro-batch (a single pair = the plain step above).  Chunk c
    covers ``X[c*n:(c+1)*n]`` of this rank's samples and, in ``indices_c``, the same slice of EVERY rank's samples
    (rank-major), i.e. the global micro-batch this rank's tables are looked up for.  MLP gradients accumulate over
    the chunks; each chunk's loss is scaled by 1/n_chunks so that their sum is the batch mean the reference
    back-propagates.  ``post_backward`` (the MLP-gradient all-reduce) runs once, before the optimizer step.
    Returns the batch-mean loss tensor."""
    X, Y = batch[0], batch[1]
    pairs = [(batch[2 + 2 * c], batch[3 + 2 * c]) for c in range((len(batch) - 2) // 2)]
    n = len(pairs)
    optimizer.zero_grad(set_to_none=True)
    if n > 1:
        if dlrm.peer_exchanges is None or len(dlrm.peer_exchanges) != n:
            raise RuntimeError("a micro-batched step needs enable_peer_exchange(dlrm, local_batch, n_chunks=%d)" % n)
        dlrm.begin_step(pairs)
    per = X.shape[0] // n
    total = None
    for c, (idx, off) in enumerate(pairs):
        Xc, Yc = (X, Y) if n == 1 else (X[c * per:(c + 1) * per], Y[c * per:(c + 1) * per])
        Z = dlrm(Xc, idx, off, None, chunk=c).float()
        E = dlrm.loss_fn(Z, Yc)
        (E if n == 1 else E / n).backward()
        total = E.detach() if total is None else total + E.detach()
    if post_backward is not None:
        post_backward()
    optimizer.step()
    dlrm.emb_l.join()  # a side-stream update joins the step (and a graph capture) here
    return total if n == 1 else total / n
