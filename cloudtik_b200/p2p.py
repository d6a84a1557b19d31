"""Fused embedding + pooled all-to-all over NVLink peer memory (table-wise sharding, one process per GPU).

The reference's ``distributed_forward`` (``P/dlrm_s_pytorch.py:342-369``) pools the local tables for the
global batch, then ``ext_dist.alltoall`` (``P/extend_distributed.py:271-333``) converts table-sharded /
global-batch into batch-sharded / all-tables, and autograd mirrors it for the gradients.  Here both
exchanges disappear into the embedding kernels:

* forward: ``b200_emb_fwd_sharded`` writes every pooled row directly into the DESTINATION rank's
  interaction input ``T[lN, F, D]`` (a symmetric-memory buffer every rank maps), so after one
  cross-rank barrier the tcgen05 interaction kernel reads its tile in place -- no send buffer, no
  NCCL all-to-all, no re-packing copy;
* backward: the interaction backward writes ``dT[lN, F, D]`` locally; after a barrier
  ``b200_emb_bwd_apply_sharded`` gathers the gradient rows it needs from the owning ranks' ``dT`` while
  it reduces and updates.

Two barriers per step order everything (kernels are stream-ordered on each rank):
  B1 after emb_fwd_sharded  : all rows of T have landed           (also: peers finished reading my dT)
  B2 after interact_bwd     : all dT are complete                 (also: nobody still needs last T)
Single box only (NVLink / NVSwitch peers); uses ``torch.distributed._symmetric_memory`` for the mapping.

Micro-batch pipeline.  Both exchange kernels are bound by NVLink (at 8 ranks 7/8 of every pooled row and of every
gradient row crosses it), and inside ONE batch the dependency chain exchange -> interaction -> top MLP -> interaction
backward -> gradient exchange leaves only the small bottom MLP to hide them behind.  A step may therefore be cut
into ``n_chunks`` micro-batches (chunk c = every rank's c-th slice of its local batch), each with its own ``T`` / ``dT``
buffers, barriers and backward workspace.  All chunks' forward exchanges are issued on the embedding's side stream at
the start of the step (they all read the tables BEFORE any update, exactly like one big batch); chunk c+1's exchange
then runs under chunk c's interaction / top MLP, and chunk c's gradient-gathering update runs under chunk c+1's
compute.  The updates are applied chunk after chunk (SGD: W -= lr*g_0, then W -= lr*g_1 -- the sum of the reference's
single update up to fp32 rounding).  See ``cloudtik_b200.graph.train_step_chunked`` for the driver of one step.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import ops


class PeerExchange:
    """Symmetric T / dT buffers of one model replica plus their per-rank pointer tables."""

    def __init__(self, lN: int, F: int, D: int, dtype: torch.dtype, device, chunk: int = 0):
        """``lN``: samples per rank in one (micro-)batch; ``chunk``: which micro-batch of the step these buffers
        serve (a step cut into n chunks owns n PeerExchange objects, see the module docstring)."""
        import torch.distributed._symmetric_memory as symm
        self.chunk = int(chunk)
        self.lN, self.F, self.D, self.dtype = lN, F, D, dtype
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        group = dist.group.WORLD.group_name
        self.T = symm.empty((lN, F, D), dtype=dtype, device=device)
        self.dT = symm.empty((lN, F, D), dtype=dtype, device=device)
        self.hT = symm.rendezvous(self.T, group)
        self.hdT = symm.rendezvous(self.dT, group)
        self.T_ptrs_dev = int(self.hT.buffer_ptrs_dev)
        self.dT_ptrs_dev = int(self.hdT.buffer_ptrs_dev)
        # every rank's dT as a tensor over peer memory (gather_grads)
        self.dT_peers = [self.hdT.get_buffer(r, (lN, F, D), dtype) for r in range(self.world)]
        self.T.zero_()
        self.dT.zero_()
        torch.cuda.synchronize()
        dist.barrier()

    def barrier(self, channel: int = 0):
        self.hT.barrier(channel=channel)

    def gather_grads(self, feat0: int, n_feat: int) -> torch.Tensor:
        """Copy the gradient rows of feature slots ``feat0 .. feat0 + n_feat`` of EVERY rank's dT into one local
        ``[B_glob, n_feat, D]`` tensor (the mirror all-to-all as W strided peer reads).  For multi-hot bags the fused
        update reads each bag's gradient row once per index: from a local copy those re-reads hit L2 / HBM, whereas
        peer memory is not cached in the local L2 and every one of them would cross NVLink again."""
        G = torch.empty(self.lN * self.world, n_feat, self.D, dtype=self.dtype, device=self.T.device)
        for k in range(self.world):  # start with the next rank: the ranks then read from different peers at a time
            r = (self.rank + 1 + k) % self.world
            G[r * self.lN:(r + 1) * self.lN].copy_(self.dT_peers[r][:, feat0:feat0 + n_feat])
        return G


class _ShardedEmbInteractFn(torch.autograd.Function):
    """x [lN, D] (+ this rank's tables, global-batch indices) -> interaction output R [lN, ldr]."""

    @staticmethod
    def forward(ctx, x, trigger, emb, exch, indices, offsets, feat0, pad_to, rowwise=None, rw_indices=None,
                rw_offsets=None, train=True, repl=None, repl_indices=None, repl_offsets=None, *repl_weights):
        B_glob = exch.lN * exch.world
        if repl is not None:
            # replicated small tables: LOCAL batch, feature slots 1 .. n right behind x.  Local reads and writes only,
            # so it goes first and runs under the tail of the exchange the next lines wait for
            if not repl.wait_pool(repl_indices, repl_offsets):  # (else: launched ahead of the bottom MLP)
                repl.pool_into(repl_indices, repl_offsets, exch.lN, exch.T[:, 1:1 + repl.n_tables], exch.D,
                               exch.F * exch.D)
                if train:
                    repl.launch_prepare(repl_indices, repl_offsets, exch.lN)
        if emb is not None:
            pend = emb._take_pending_sharded(exch.chunk)
            if pend is not None and pend[0] is indices and pend[1] is offsets:
                torch.cuda.current_stream().wait_event(pend[2])  # launched ahead of the bottom MLP (prelaunch())
            else:
                emb.join()
                with emb._timed("emb_fwd"):
                    ops.emb_fwd_sharded(emb.tables, indices, offsets, B_glob, exch.T_ptrs_dev, exch.lN, exch.F,
                                        feat0, exch.dtype, emb.error_flag, first_dest=(exch.rank + 1) % exch.world)
                if train:  # sort / dedup on the side stream: only a backward consumes it
                    emb._launch_prepare(indices, offsets, B_glob, chunk=exch.chunk)
        ctx.rw_ctx = None
        if rowwise is not None:
            # row-wise sharded hot tables occupy the LAST feature slots: every rank ADDS its partial pooled rows
            # into every rank's T over peer memory (each rank zeroed its slots before the previous step's B2, or
            # when the buffers were created)
            ctx.rw_ctx = rowwise.pool_to_peers(rw_indices, rw_offsets, exch)
        exch.T[:, 0].copy_(x)
        exch.barrier(0)  # B1
        R = ops.interact_fwd(exch.T, pad_to=pad_to)
        if not train:
            # forward-only (eval / no_grad): there is no backward and hence no B2 -- without a second barrier a
            # faster rank could start the NEXT forward and store into this rank's T while interact_fwd above is
            # still reading it
            if rowwise is not None:
                exch.T[:, exch.F - rowwise.k:].zero_()
            exch.barrier(1)
        ctx.emb, ctx.exch, ctx.feat0, ctx.rowwise = emb, exch, feat0, rowwise
        ctx.repl, ctx.repl_bags = repl, (repl_indices, repl_offsets)
        ctx.save_for_backward(indices, offsets)
        ctx.x_dtype = x.dtype
        return R

    @staticmethod
    def backward(ctx, dR):
        emb, exch = ctx.emb, ctx.exch
        indices, offsets = ctx.saved_tensors
        B_glob = exch.lN * exch.world
        ops.interact_bwd(exch.T, dR.contiguous().to(exch.dtype), out=exch.dT)
        if ctx.rowwise is not None:
            # T is read for the last time above: clear the row-wise slots for the next step's partial sums.  A peer
            # starts adding into them only after it passed B2, i.e. after this
            exch.T[:, exch.F - ctx.rowwise.k:].zero_()
        exch.barrier(1)  # B2
        if emb is not None:
            emb._apply_update_sharded(indices, offsets, B_glob, exch, ctx.feat0)
        if ctx.rowwise is not None:
            ctx.rowwise.update_from_peers(ctx.rw_ctx, exch)
        dx = exch.dT[:, 0].clone()  # dT is overwritten by the next step; the bottom MLP backward runs now
        rg = ()
        if ctx.repl is not None:  # dense gradients of the replicated tables (all-reduced with the MLPs' by the caller)
            rg = tuple(ctx.repl.dense_grads(ctx.repl_bags[0], ctx.repl_bags[1], exch.lN,
                                            exch.dT[:, 1:1 + ctx.repl.n_tables], exch.D, exch.F * exch.D))
        return (dx.to(ctx.x_dtype),) + (None,) * 14 + rg


def prelaunch(emb, exch, indices, offsets, feat0: int):
    """Start this rank's pooled forward + exchange on the embedding's side stream before the caller runs the
    bottom MLP; ``sharded_emb_interact`` with the same (indices, offsets) tensors picks it up.
    Micro-batched step: pass lists (one PeerExchange / indices / offsets per chunk) -- every chunk's exchange is
    queued now, in chunk order, followed by every chunk's index stage."""
    if emb is None or not (emb.overlap_sharded and emb.overlap_prepare and emb.profile is None):
        return
    if isinstance(exch, PeerExchange):
        exch, indices, offsets = [exch], [indices], [offsets]
    emb._launch_forward_sharded_async(list(zip(indices, offsets, exch)), feat0)


def sharded_emb_interact(x, emb, exch: PeerExchange, indices, offsets, feat0: int, pad_to: int = 1, rowwise=None,
                         rw_indices=None, rw_offsets=None, trigger=None, train=None, repl=None, repl_indices=None,
                         repl_offsets=None):
    """``train`` (default: grad mode is on): False = forward only -- no index stage is launched for a backward that
    will not come, and a second barrier replaces the one the backward would have issued.
    ``repl``: a ``ReplicatedEmbeddingBags`` whose tables fill the feature slots right behind x (``feat0`` of the
    sharded tables then starts behind them), looked up for the LOCAL bags ``repl_indices`` / ``repl_offsets``."""
    trig = trigger if trigger is not None else (emb._trigger if emb is not None else rowwise.emb._trigger)
    if train is None:
        train = torch.is_grad_enabled()
    rw = tuple(repl.params()) if repl is not None else ()
    return _ShardedEmbInteractFn.apply(x, trig, emb, exch, indices, offsets, feat0, pad_to, rowwise, rw_indices,
                                       rw_offsets, bool(train), repl, repl_indices, repl_offsets, *rw)


def split_global_batch(idx_l, off_l, world: int, n_chunks: int):
    """Cut per-table bags of a GLOBAL batch (bag b belongs to rank ``b // lN``) into ``n_chunks`` micro-batches:
    micro-batch c holds, rank after rank, the c-th slice of every rank's samples -- the bag order
    ``b200_emb_fwd_sharded`` expects for a chunk.  ``idx_l[t]`` int64[nnz_t], ``off_l[t]`` int64[B_glob] start
    offsets.  Returns a list over chunks of (idx_l_c, off_l_c).  Host-side helper (tests, the training CLI): it
    synchronises; a production input pipeline emits its batches in this order in the first place."""
    out = [([], []) for _ in range(n_chunks)]
    for idx, off in zip(idx_l, off_l):
        B = off.numel()
        lN = B // world
        per = lN // n_chunks
        ends = torch.cat([off[1:], torch.tensor([idx.numel()], dtype=off.dtype, device=off.device)])
        lens = ends - off
        bag = torch.arange(B, device=off.device)
        chunk_of_bag = (bag % lN) // per
        elem_chunk = torch.repeat_interleave(chunk_of_bag, lens)
        for c in range(n_chunks):
            lc = lens[chunk_of_bag == c]
            oc = torch.cumsum(lc, 0) - lc
            out[c][0].append(idx[elem_chunk == c])
            out[c][1].append(oc)
    return out
