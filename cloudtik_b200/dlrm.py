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
        ids = range(ln.size) if local_ln_emb is None else local_ln_emb
        weights = self._draw_tables(m, ln, ids, np_init_emb_weight, lambda i: device)
        return MergedEmbeddingBagWithSGD(weights, lr=lr, optimizer=optimizer)

    def _draw_tables(self, m, ln, ids, np_init_emb_weight, device_of):
        """Tables ``ids`` drawn in that order (the order fixes the numpy stream), table i placed on ``device_of(i)``."""
        weights = []
        for i in ids:
            n = int(ln[i])
            if np_init_emb_weight:
                ra = np.random.RandomState(self.l_emb_seeds[i]) if self.np_emb_seeds else np.random
                W = ra.uniform(low=-np.sqrt(1 / n), high=np.sqrt(1 / n), size=(n, m)).astype(np.float32)
                weights.append(torch.from_numpy(W).to(device_of(i)))
            else:
                w = torch.empty(n, m, dtype=torch.float32, device=device_of(i))
                w.uniform_(-float(np.sqrt(1 / n)), float(np.sqrt(1 / n)))
                weights.append(w)
        return weights

    def __init__(self, m_spa=None, ln_emb=None, ln_bot=None, ln_top=None, sigmoid_bot=-1, sigmoid_top=-1,
                 weighted_pooling=None, loss_threshold=0.0, np_init_emb_weight=False, device="cuda",
                 lr=0.01, emb_optimizer="sgd", bf16=False, n_rowwise=0, distributed=None,
                 sparse_dense_boundary=0, dense_tables_first=False, arch_interaction_op="dot",
                 arch_interaction_itself=False, np_emb_seeds=False, ndevices=-1):
        """``distributed=False`` builds the single-process model (all tables local) even inside an initialised
        process group -- the oracle of the multi-GPU parity checks runs that way next to the sharded model.

        ``sparse_dense_boundary`` (T/dlrm_s_pytorch.py:266-269,353-358; 0 = off, the product variant's behaviour):
        when distributed, tables with fewer rows are REPLICATED on every rank (``emb_dense``: dense parameters,
        local-batch lookups, gradients all-reduced with the MLPs') and only the others are sharded table-wise; the
        interaction then sees ``[x] + dense tables + sharded tables`` -- the reference's permuted feature order
        (``ly = ly_dense + list(ly_sparse)``, T/dlrm_s_pytorch.py:509).  ``dense_tables_first=True`` gives the
        single-process model the same table order (``table_order``), which makes it the parity oracle of such a
        run (the reference's own sequential_forward keeps the natural order, SURVEY.md section 8c caveat)."""
        super().__init__()
        # single-process multi-GPU mode of the training variant (T/dlrm_s_pytorch.py:306,524-630): see parallel_forward
        self.ndevices = int(ndevices)
        # model options of the training variant (T/dlrm_s_pytorch.py:289-305,407-444)
        if arch_interaction_op not in ("dot", "cat"):
            sys.exit("ERROR: --arch-interaction-op=" + str(arch_interaction_op) + " is not supported")
        self.arch_interaction_op = arch_interaction_op
        self.arch_interaction_itself = bool(arch_interaction_itself)
        # np_emb_seeds: every table is drawn from its own RandomState(l_emb_seeds[i]) like the training variant
        # (T/dlrm_s_pytorch.py:240-243) instead of the global stream after the MLPs (product variant)
        self.np_emb_seeds = bool(np_emb_seeds)
        self.world = ext_dist.my_size if (distributed is None or distributed) and ext_dist.my_size > 1 else 1
        if self.world > 1 and (self.arch_interaction_itself or self.arch_interaction_op != "dot"):
            raise NotImplementedError("the sharded paths implement the plain dot interaction only")
        if weighted_pooling is not None:
            raise NotImplementedError("weighted pooling is not on the reference's product path")
        self.loss_threshold = loss_threshold
        self.bf16 = bf16
        self.fused_mlp = True
        self.peer_exchange = None  # set by enable_peer_exchange(): fused a2a over NVLink peer memory
        self.peer_exchanges = None  # one per micro-batch of a step (peer_exchange is the first)
        self.exchange_first = None  # peer-memory path: start the exchange before the bottom MLP (None = by world size)
        ln_emb = np.asarray(ln_emb)
        # the LAST n_rowwise tables are row-wise sharded over the ranks (hot tables, config 5); the others
        # are partitioned table-wise exactly like the reference (P/dlrm_s_pytorch.py:256-264)
        self.n_rowwise = int(n_rowwise) if self.world > 1 else 0
        self.rowwise = None
        self.emb_dense = None
        self.sparse_dense_boundary = int(sparse_dense_boundary)
        n_tw = len(ln_emb) - self.n_rowwise
        split = self.sparse_dense_boundary > 0 and (self.world > 1 or dense_tables_first)
        self.ln_emb_dense = [i for i in range(n_tw) if split and ln_emb[i] < self.sparse_dense_boundary]
        self.ln_emb_sparse = [i for i in range(n_tw) if i not in self.ln_emb_dense]
        # order of the tables as interaction features (identity unless small tables are moved to the front)
        self.table_order = self.ln_emb_dense + self.ln_emb_sparse + list(range(n_tw, len(ln_emb)))
        if self.world > 1:
            n_emb = len(self.ln_emb_sparse)
            self.n_global_emb = len(ln_emb)
            self.rank = ext_dist.dist.get_rank()
            self.ln_emb = list(self.ln_emb_sparse)
            self.n_local_emb, self.n_emb_per_rank = ext_dist.get_split_lengths(n_emb)
            self.local_ln_emb_slice = ext_dist.get_my_slice(n_emb)
            self.local_ln_emb = self.ln_emb[self.local_ln_emb_slice]
        else:
            self.local_ln_emb = None if not self.ln_emb_dense else list(self.table_order)
        self.l_emb_seeds = np.random.randint(low=0, high=100000, size=len(ln_emb))
        self.bot_l = self.create_mlp(np.asarray(ln_bot), sigmoid_bot, pad_first=True).to(device)
        self.dense_pad = (-int(ln_bot[0])) % 8  # 13 dense features -> 16 zero-padded columns (aligned GEMM K)
        self.top_l = self.create_mlp(np.asarray(ln_top), sigmoid_top, pad_first=True).to(device)
        self.interaction_pad = 8  # interaction output rounded up to a multiple of 8 columns
        if self.ndevices > 1:
            if self.world > 1 or self.ln_emb_dense or self.n_rowwise:
                raise NotImplementedError("ndevices > 1 is the single-process mode: no process group, no table split")
            if str(torch.device(device)) not in ("cuda", "cuda:0"):
                raise RuntimeError("ndevices > 1 keeps the MLP parameters on cuda:0 like the reference")
            # (the reference's caller passes min(#GPUs, batch, #tables), T/dlrm_s_pytorch.py:938)
            self.ndevices = min(self.ndevices, len(ln_emb), max(torch.cuda.device_count(), 1))
            if self.ndevices < 2:
                raise RuntimeError("ndevices > 1 needs at least two visible GPUs and two tables")
            self.device_ids = list(range(self.ndevices))
            # table k lives on device k % ndevices (T/dlrm_s_pytorch.py:543-548), one merged module per device
            ws = self._draw_tables(m_spa, ln_emb, range(len(ln_emb)), np_init_emb_weight,
                                   lambda i: torch.device("cuda", i % self.ndevices))
            self.tables_of_device = [list(range(d, len(ln_emb), self.ndevices)) for d in self.device_ids]
            self.emb_l = nn.ModuleList(
                [MergedEmbeddingBagWithSGD([ws[k] for k in ks], lr=lr, optimizer=emb_optimizer)
                 for ks in self.tables_of_device])
            for mod in self.emb_l:
                mod.reserve_dense_slot = False  # their pooled rows travel before they reach an interaction input
                if bf16:
                    mod.output_dtype = torch.bfloat16
            self.need_linearize_indices_and_offsets = torch.BoolTensor([True])
            self.loss_fn = torch.nn.BCELoss(reduction="mean")
            return
        self.emb_l = self.create_emb(m_spa, ln_emb, self.local_ln_emb, np_init_emb_weight, device, lr,
                                     emb_optimizer)
        if self.world > 1 and self.ln_emb_dense:
            from .replicated import ReplicatedEmbeddingBags
            small = self.create_emb(m_spa, ln_emb, self.ln_emb_dense, np_init_emb_weight, device, lr, "sgd")
            self.emb_dense = ReplicatedEmbeddingBags(small._raw_weights())
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
        if self.world == 1 and self.ln_emb_dense:  # natural table order in, dense-first order inside
            lS_o, lS_i = [lS_o[t] for t in self.table_order], [lS_i[t] for t in self.table_order]
        return emb_l((list(lS_i), list(lS_o), [False] * len(lS_i)), torch.BoolTensor([True]))

    def interact_features(self, x, ly):
        if self.arch_interaction_op == "cat":
            from .interaction import interaction_cat
            return interaction_cat(x, *ly)
        return interaction(x, *ly, pad_to=self.interaction_pad, itself=self.arch_interaction_itself)

    def forward(self, dense_x, *emb_args, chunk=0, dense_bags=None):
        """``chunk``: which micro-batch of a step cut by ``enable_peer_exchange(..., n_chunks)`` this call is
        (0 when the step is not cut).  ``dense_bags``: (indices, offsets) of this rank's LOCAL bags of the replicated
        small tables, linearised in ``ln_emb_dense`` order (models built with ``sparse_dense_boundary`` only)."""
        if self.world > 1:
            return self.distributed_forward(dense_x, *emb_args, chunk=chunk, dense_bags=dense_bags)
        if self.ndevices > 1:
            return self.parallel_forward(dense_x, *emb_args)
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
                  1 + len(self.ln_emb_dense) + self.local_ln_emb_slice.start)

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

    def parallel_forward(self, dense_x, lS_o, lS_i):
        """One process driving ``ndevices`` GPUs (T/dlrm_s_pytorch.py:524-630): MLPs data-parallel (replicas of the
        cuda:0 parameters, re-made every call so their gradients sum back onto cuda:0), tables model-parallel (table k
        on device k % ndevices, each device pools its tables for the WHOLE batch with one ``b200_emb_fwd`` launch), then
        every device receives its batch slice of every table (the reference's "butterfly shuffle": here one
        device-to-device copy per (source, destination) pair instead of one per table), interacts and runs the top MLP
        on its slice; the predictions are gathered on cuda:0.  One host thread issues everything: the launches are
        asynchronous, so the devices run concurrently without ``parallel_apply``'s worker threads.  The fused sparse
        update of a device's tables runs inside backward on that device, with the full learning rate (the loss is the
        mean over the gathered batch).  The device count is fixed at construction: a batch smaller than it is refused
        (the reference re-distributes the model instead)."""
        from torch.nn.parallel import gather, replicate, scatter
        devs, nd = self.device_ids, len(self.device_ids)
        batch_size = dense_x.size()[0]
        if batch_size < nd:
            sys.exit("ERROR: batch_size (%d) must be at least the number of devices (%d) in parallel_forward call"
                     % (batch_size, nd))
        n_tables = sum(len(ks) for ks in self.tables_of_device)
        if len(lS_o) != n_tables or len(lS_i) != n_tables:
            sys.exit("ERROR: corrupted model input detected in parallel_forward call")
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.bf16):
            bot = replicate(self.bot_l, devs)
            top = replicate(self.top_l, devs)
            xs = scatter(self._pad_dense(dense_x), devs, dim=0)
            pooled = []  # per source device: [B, T_d, D] pooled rows of its tables, whole batch
            for d, ks in enumerate(self.tables_of_device):
                dev = torch.device("cuda", devs[d])
                with torch.cuda.device(dev):
                    ly = self.emb_l[d](([lS_i[k].to(dev, non_blocking=True) for k in ks],
                                        [lS_o[k].to(dev, non_blocking=True) for k in ks], [False] * len(ks)), True)
                    pooled.append(torch.stack(list(ly), dim=1))
            x = []
            for d in range(nd):
                with torch.cuda.device(devs[d]):
                    x.append(bot[d](xs[d]))
            # slices[s][j]: rows of device j's batch slice from source device s (same cut as the dense features)
            slices = [scatter(t, devs, dim=0) for t in pooled]
            p = []
            for j in range(nd):
                with torch.cuda.device(devs[j]):
                    ly_j = [slices[k % nd][j][:, k // nd] for k in range(n_tables)]
                    p.append(top[j](self.interact_features(x[j], ly_j)))
            p0 = gather(p, devs[0], dim=0)
        return self._clamp(p0)

    def distributed_forward(self, dense_x, *emb_args, chunk=0, dense_bags=None):
        # P/dlrm_s_pytorch.py:342-369: local tables on the GLOBAL batch -> all-to-all -> local batch
        batch_size = dense_x.size()[0]
        if batch_size < ext_dist.my_size:
            sys.exit("ERROR: batch_size (%d) must be larger than number of ranks (%d)"
                     % (batch_size, ext_dist.my_size))
        if self.peer_exchange is not None:
            return self._p2p_forward(dense_x, *emb_args, chunk=chunk, dense_bags=dense_bags)
        if self.emb_dense is not None:
            raise NotImplementedError("replicated small tables (sparse_dense_boundary) need the peer-memory "
                                      "exchange: call enable_peer_exchange(), or build the model with boundary 0")
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


def _p2p_forward(self, dense_x, *emb_args, chunk=0, dense_bags=None):
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
        if len(lS_i) == self.n_global_emb - self.n_rowwise and len(lS_i) != len(self.local_ln_emb):
            # the training variant's input (T/dlrm_s_pytorch.py:489-497): EVERY table's bags for this rank's LOCAL
            # batch.  Small tables stay here; the sharded tables' bags travel to their owners (index all-to-all)
            if self.emb_dense is not None:
                dense_bags = self.emb_dense_linearize([lS_i[t] for t in self.ln_emb_dense],
                                                      [lS_o[t] for t in self.ln_emb_dense])
            lS_o, lS_i = ext_dist.shuffle_bags([lS_o[t] for t in self.ln_emb_sparse],
                                               [lS_i[t] for t in self.ln_emb_sparse], self.n_emb_per_rank)
        indices, offsets, _ = self.emb_l.linearize_indices_and_offsets(list(lS_i), list(lS_o), [False] * len(lS_i))
    if (self.emb_dense is not None) != (dense_bags is not None):
        raise RuntimeError("replicated small tables need dense_bags=(indices, offsets) of the local batch")
    if (self.rowwise is not None) != (rw_i is not None):
        raise RuntimeError("row-wise sharded tables need (indices, offsets, None, rw_indices, rw_offsets)")
    feat0 = 1 + len(self.ln_emb_dense) + self.local_ln_emb_slice.start
    train = self.training and torch.is_grad_enabled()
    # When does the NVLink-bound exchange start on the side stream?  Launched ahead of the (short) bottom MLP, its
    # thousands of long-lived thread blocks hold every SM slot and stretch the MLP's small kernels to about 150 us;
    # that only pays when the exchange itself is the longer of the two, i.e. from 4 ranks on (230 us at 8 ranks,
    # 105 us at 2).  Either way the replicated tables' local lookup runs next to it.
    first = train and len(self.peer_exchanges) == 1 and (self.exchange_first if self.exchange_first is not None
                                                          else exch.world > 2)
    if first:
        prelaunch(self.emb_l, exch, indices, offsets, feat0)
    if train and self.emb_dense is not None and len(self.peer_exchanges) == 1:
        # the replicated tables' local lookup (and the index stage of their backward) next to the bottom MLP too
        self.emb_dense.pool_into_async(dense_bags[0], dense_bags[1], exch.lN, exch.T[:, 1:1 + self.emb_dense.n_tables],
                                       exch.D, exch.F * exch.D)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.bf16):
        x = self.apply_mlp(self._pad_dense(dense_x), self.bot_l)
        if train and len(self.peer_exchanges) == 1 and not first:
            prelaunch(self.emb_l, exch, indices, offsets, feat0)
        z = sharded_emb_interact(x, self.emb_l, exch, indices, offsets, feat0,
                                 pad_to=self.interaction_pad, rowwise=self.rowwise, rw_indices=rw_i,
                                 rw_offsets=rw_o, train=train, repl=self.emb_dense,
                                 repl_indices=None if dense_bags is None else dense_bags[0],
                                 repl_offsets=None if dense_bags is None else dense_bags[1])
        p = self.apply_mlp(z, self.top_l)
    return self._clamp(p)


def _emb_dense_linearize(self, idx_l, off_l):
    """Per-table LOCAL bags of the replicated tables -> (indices, offsets) in the merged layout."""
    dev = idx_l[0].device
    nnz, parts = 0, []
    for i, o in zip(idx_l, off_l):
        parts.append(o.to(torch.int64) + nnz)
        nnz += i.numel()
    parts.append(torch.tensor([nnz], dtype=torch.int64, device=dev))
    return torch.cat([i.to(torch.int64) for i in idx_l]), torch.cat(parts)


DLRM_Net._p2p_forward = _p2p_forward
DLRM_Net.emb_dense_linearize = _emb_dense_linearize


def enable_peer_exchange(dlrm, local_batch: int, n_chunks: int = 1):
    """Switch a table-wise sharded DLRM_Net (one process per GPU on one NVLink box) to the fused
    peer-memory exchange; ``local_batch`` = samples per rank (fixed: the buffers are symmetric).
    ``n_chunks`` > 1 cuts every step into that many micro-batches (``local_batch`` must divide), each with its own
    exchange buffers, so that one chunk's NVLink traffic runs under another chunk's compute (cloudtik_b200.p2p)."""
    from .p2p import PeerExchange
    if n_chunks < 1 or local_batch % n_chunks:
        raise ValueError("enable_peer_exchange: the local batch must be a multiple of n_chunks")
    if n_chunks > 1 and (dlrm.rowwise is not None or dlrm.emb_dense is not None):
        raise NotImplementedError("micro-batched steps with row-wise sharded or replicated tables")
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


def train_step_chunked(dlrm, optimizer, batch, post_backward=None, loss_scale=1.0):
    """One iteration on a batch given as the flat tuple ``(X, Y, indices_0, offsets_0[, indices_1, offsets_1, ...])``:
    one (indices, offsets) pair of linearised bags per micro-batch (a single pair = the plain step above), followed --
    only for a model with replicated small tables (``emb_dense``) -- by one more pair: this rank's LOCAL bags of
    those tables.  Chunk c
    covers ``X[c*n:(c+1)*n]`` of this rank's samples and, in ``indices_c``, the same slice of EVERY rank's samples
    (rank-major), i.e. the global micro-batch this rank's tables are looked up for.  MLP gradients accumulate over
    the chunks; each chunk's loss is scaled by 1/n_chunks so that their sum is the batch mean the reference
    back-propagates (``loss_scale``: one more factor on the back-propagated loss, e.g. 1/world when a scale-sensitive
    table optimiser wants the MLP gradients summed instead of averaged).  ``post_backward`` (the MLP-gradient all-reduce) runs once, before the optimizer step.
    Returns the batch-mean loss tensor."""
    X, Y = batch[0], batch[1]
    pairs = [(batch[2 + 2 * c], batch[3 + 2 * c]) for c in range((len(batch) - 2) // 2)]
    # trailing pairs: GLOBAL-batch bags of the row-wise sharded tables (global row ids), and before them this rank's
    # LOCAL bags of the replicated small tables
    rw_bags = pairs.pop() if getattr(dlrm, "rowwise", None) is not None else None
    dense_bags = pairs.pop() if getattr(dlrm, "emb_dense", None) is not None else None
    n = len(pairs)
    optimizer.zero_grad(set_to_none=True)
    if hasattr(post_backward, "begin"):
        post_backward.begin(n)
    if n > 1:
        if dlrm.peer_exchanges is None or len(dlrm.peer_exchanges) != n:
            raise RuntimeError("a micro-batched step needs enable_peer_exchange(dlrm, local_batch, n_chunks=%d)" % n)
        dlrm.begin_step(pairs)
    per = X.shape[0] // n
    total = None
    rep = getattr(dlrm, "emb_dense", None)
    if rep is not None:
        rep.defer = True  # dense gradients of the replicated tables off the main chain; joined before the all-reduce
    for c, (idx, off) in enumerate(pairs):
        Xc, Yc = (X, Y) if n == 1 else (X[c * per:(c + 1) * per], Y[c * per:(c + 1) * per])
        emb_args = (idx, off, None) if rw_bags is None else (idx, off, None, rw_bags[0], rw_bags[1])
        Z = (dlrm(Xc, *emb_args, chunk=c) if dense_bags is None
             else dlrm(Xc, *emb_args, chunk=c, dense_bags=dense_bags)).float()
        E = dlrm.loss_fn(Z, Yc)
        (E * (loss_scale / n) if (n > 1 or loss_scale != 1.0) else E).backward()
        total = E.detach() if total is None else total + E.detach()
    if rep is not None:
        rep.defer = False
        rep.join()
    if post_backward is not None:
        post_backward()
    optimizer.step()
    dlrm.emb_l.join()  # a side-stream update joins the step (and a graph capture) here
    return total if n == 1 else total / n
