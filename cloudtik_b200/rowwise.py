"""Row-wise sharded hot tables (BASELINE config 5; SURVEY.md section 8e -- not in the reference, which only
shards table-wise).  Rank r of W owns rows {i : i mod W == r} of every hot table, stored densely as local
row i // W.  For the GLOBAL batch every rank pools the rows it owns (partial sums); the partial sums of the W ranks
have to be added up for each sample's rank (a reduce-scatter over the batch dimension), and in backward every rank
needs the output gradients of ALL samples (an all-gather) to run the fused sort / dedup / update on its rows.

Both exchanges are fused into the embedding kernels over NVLink peer memory, like the table-wise ones
(``cloudtik_b200.p2p``):
* ``b200_rowwise_filter`` keeps the owned indices of every bag in a FIXED-capacity array (fill level on the device:
  no data-dependent shape, no host sync -- the step is graph-capturable);
* forward: ``b200_emb_fwd_sharded_add`` ADDS each partial pooled row straight into the destination rank's slot of
  the interaction input (``red.global.add`` over peer memory; the destination zeroed the slot; bags without an
  owned index send nothing) -- the reduce-scatter;
* backward: ``b200_emb_bwd_prepare_partial`` + ``b200_emb_bwd_apply_sharded`` read the gradient rows of every
  rank's samples from the owning rank's ``dT`` -- the all-gather.
(``pool`` / ``update`` keep the NCCL reduce_scatter / all_gather formulation: the reference point the fused one was
checked against, and the path for a job without peer access.)
Sharded-table learning rate is lr / W for the same reason as table-wise sharding (every rank contributes the
gradient of a local-mean loss).
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn

from . import ops
from .merged_embedding import MergedEmbeddingBagWithSGD


def filter_owned(indices, offsets, world: int, rank: int):
    """CSR bags over global row ids -> CSR bags over the rows this rank owns (local ids = id // world)."""
    mask = (indices % world) == rank
    csum = torch.zeros(indices.numel() + 1, dtype=torch.int64, device=indices.device)
    torch.cumsum(mask, 0, out=csum[1:])
    return torch.div(indices[mask], world, rounding_mode="floor"), csum[offsets]


class RowWiseShardedTables(nn.Module):
    def __init__(self, full_weights_or_rows, D, world, rank, device, lr=0.01, optimizer="sgd", out_dtype=None):
        """``full_weights_or_rows``: per hot table either its full fp32 weight (host or device tensor; the owned
        rows are sliced out) or just its row count (owned rows are drawn U(-sqrt(1/n), sqrt(1/n)) on device)."""
        super().__init__()
        self.world, self.rank = world, rank
        shards, self.full_rows = [], []
        for w in full_weights_or_rows:
            if torch.is_tensor(w):
                self.full_rows.append(int(w.shape[0]))
                shards.append(w[rank::world].to(device=device, dtype=torch.float32).contiguous())
            else:
                n = int(w)
                self.full_rows.append(n)
                local = (n - rank + world - 1) // world
                s = torch.empty(max(local, 1), D, dtype=torch.float32, device=device)
                s.uniform_(-(1.0 / n) ** 0.5, (1.0 / n) ** 0.5)
                shards.append(s)
        self.k = len(shards)
        self.dim = D
        self.emb = MergedEmbeddingBagWithSGD(shards, lr=lr, optimizer=optimizer)
        self.emb.reserve_dense_slot = False
        self.emb.output_dtype = out_dtype
        self.emb.overlap_prepare = False
        self.error_flag = None  # int32[1]: 2 = a step owned more indices than the fixed capacity (bags truncated)

    def pool(self, indices, offsets, lN: int):
        """indices / offsets: linearised GLOBAL-batch bags of the k hot tables (global row ids).
        Returns ([lN, k, D] pooled rows of this rank's samples, ctx for ``update``)."""
        B_glob = lN * self.world
        li, lo = filter_owned(indices, offsets, self.world, self.rank)
        dt = self.emb.output_dtype or self.emb.tables.dtype
        partial = torch.empty(B_glob, self.k, self.dim, dtype=dt, device=indices.device)
        ops.emb_fwd(self.emb.tables, li, lo, B_glob, partial, self.dim, self.k * self.dim, self.emb.error_flag)
        out = torch.empty(lN, self.k, self.dim, dtype=dt, device=indices.device)
        dist.reduce_scatter_tensor(out, partial)
        return out, (li, lo, B_glob)

    def update(self, grad_local, ctx):
        """grad_local [lN, k, D] (gradient w.r.t. ``pool``'s output) -> fused update of the owned rows."""
        li, lo, B_glob = ctx
        g = grad_local.contiguous()
        gall = torch.empty(B_glob, self.k, self.dim, dtype=g.dtype, device=g.device)
        dist.all_gather_into_tensor(gall, g)
        self.emb._apply_update(li, lo, B_glob, gall, self.dim, self.k * self.dim)

    # ---- fused over peer memory -------------------------------------------------------
    def capacity(self, nnz: int) -> int:
        """Fixed size of the filtered index array: the expected share nnz / W with 25 % head-room (row ids spread
        by ``mod W``; a fuller step sets the error flag instead of reallocating)."""
        return max(1024, int(nnz / self.world * 1.25) + 1024) // 2 * 2

    def pool_to_peers(self, indices, offsets, exch):
        """indices / offsets: linearised GLOBAL-batch bags of the k hot tables (global row ids).  Adds this rank's
        partial pooled rows into the LAST k feature slots of every rank's ``exch.T`` (which that rank zeroed).
        Returns the ctx for ``update_from_peers``."""
        B_glob = exch.lN * self.world
        if self.error_flag is None:
            self.error_flag = torch.zeros(1, dtype=torch.int32, device=indices.device)
        li, lo = ops.rowwise_filter(indices, offsets, self.world, self.rank, self.capacity(indices.numel()),
                                    self.error_flag)
        ops.emb_fwd_sharded(self.emb.tables, li, lo, B_glob, exch.T_ptrs_dev, exch.lN, exch.F, exch.F - self.k,
                            exch.dtype, self.emb.error_flag, accumulate=True, first_dest=(exch.rank + 1) % exch.world)
        return li, lo, B_glob

    def update_from_peers(self, ctx, exch):
        """Fused update of the owned rows; the gradient rows come from every rank's ``exch.dT`` (after barrier B2)."""
        li, lo, B_glob = ctx
        emb = self.emb
        ws = emb._workspace(li.numel())
        ops.emb_bwd_prepare(emb.tables, li, lo, B_glob, ws, emb.error_flag, nnz_valid=lo[-1:])
        opt = ops.OPT_ADAGRAD if emb.optimizer_kind == "adagrad" else ops.OPT_SGD
        if li.numel() >= 2 * self.k * B_glob:
            # multi-hot: every bag's gradient row is read once per owned index -- from a local copy, not W times
            # over NVLink (peer memory is not cached in the local L2)
            G = exch.gather_grads(exch.F - self.k, self.k)
            ops.emb_bwd_apply(emb.tables, li.numel(), B_glob, G, self.dim, self.k * self.dim, emb.sgd_args.lr, ws,
                              optimizer=opt, eps=emb.eps, lr_tensor=emb.lr_tensor)
            return
        ops.emb_bwd_apply_sharded(emb.tables, li.numel(), B_glob, exch.dT_ptrs_dev, exch.lN, exch.F, exch.F - self.k,
                                  exch.dtype, emb.sgd_args.lr, ws, optimizer=opt, eps=emb.eps,
                                  lr_tensor=emb.lr_tensor)

    def full_table(self, t: int) -> torch.Tensor:
        """Reassemble hot table t from every rank's shard (tests / checkpoints)."""
        n = self.full_rows[t]
        mine = self.emb.master_weights()[t]
        per = (n + self.world - 1) // self.world
        pad = torch.zeros(per, self.dim, dtype=torch.float32, device=mine.device)
        pad[:mine.shape[0]] = mine
        allp = torch.empty(self.world, per, self.dim, dtype=torch.float32, device=mine.device)
        dist.all_gather_into_tensor(allp, pad)
        return allp.permute(1, 0, 2).reshape(-1, self.dim)[:n]  # row i = shard (i % W), local i // W
