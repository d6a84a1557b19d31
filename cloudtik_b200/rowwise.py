"""Row-wise sharded hot tables (BASELINE config 5; SURVEY.md section 8e -- not in the reference, which only
shards table-wise).  Rank r of W owns rows {i : i mod W == r} of every hot table, stored densely as local
row i // W.  For the GLOBAL batch every rank pools the rows it owns (partial sums), a reduce-scatter over
the batch dimension gives each rank the complete pooled rows of ITS samples; backward all-gathers the
output gradients and every rank runs the fused sort / dedup / update on the rows it owns.

Kernels: the same ``b200_emb_fwd`` / ``b200_emb_bwd_*`` entry points, fed with the bags filtered down to
the owned rows (a mask + prefix sum over the CSR arrays).  Collectives: NCCL reduce_scatter / all_gather
(a real exchange step: partial sums have to be added).  Sharded-table learning rate is lr / W for the same
reason as table-wise sharding (every rank contributes the gradient of a local-mean loss).
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
