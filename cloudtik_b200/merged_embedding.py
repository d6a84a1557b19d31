"""Merged sum-pooled EmbeddingBag with the optimiser step fused into backward.

Host-side mirror of the seam the reference already has for its own fast path:
``ipex.nn.modules.MergedEmbeddingBagWithSGD`` as used at reference
``P/dlrm_s_pytorch.py:288-289`` (call), ``:801`` (``from_embeddingbag_list``), ``:908-915`` /
``:1008-1014`` (``linearize_indices_and_offsets``), ``:1027`` (``sgd_args._replace(lr=)``),
``:344,357`` (``.weights`` / ``.n_tables``), ``:931`` (``to_bfloat16_train``).
Same names, same argument meaning; the arithmetic runs in libb200dlrm.so (no CPU path).

Layout: the pooled output of all tables is written by ONE kernel launch straight into a
``[B, T+1, D]`` buffer whose slot 0 is left free for the bottom-MLP output, so that the
interaction kernel (``cloudtik_b200.interaction``) reads it without the reference's
``torch.cat`` copy.  The returned list holds ``[B, D]`` views of that buffer.
"""
from __future__ import annotations

import contextlib
import weakref
from collections import namedtuple
from typing import List, Optional, Sequence

import torch
import torch.nn as nn

from . import ops

SGDArgs = namedtuple("SGDArgs", ["bf16_trail", "weight_decay", "lr"])
AdagradArgs = namedtuple("AdagradArgs", ["lr", "eps"])

# storage data_ptr -> weakref(pooled [B, F, D] buffer); lets interaction() recognise our layout
_POOLED_BUFFERS = {}


def _register_pooled(buf: torch.Tensor):
    key = buf.untyped_storage().data_ptr()
    _POOLED_BUFFERS[key] = weakref.ref(buf, lambda _r, k=key: _POOLED_BUFFERS.pop(k, None))


def lookup_pooled(t: torch.Tensor) -> Optional[torch.Tensor]:
    r = _POOLED_BUFFERS.get(t.untyped_storage().data_ptr())
    return r() if r is not None else None


class _TableHolder(nn.Module):
    """Keeps ``emb_l.{k}.weight`` state_dict keys (checkpoint contract, T/dlrm_s_pytorch.py:1415)."""

    def __init__(self, weight: torch.Tensor):
        super().__init__()
        self.weight = nn.Parameter(weight, requires_grad=False)  # updated inside backward


def _common_layout(grads: Sequence[torch.Tensor], B: int, D: int):
    """If the per-table gradients are [B, D] views at a constant element distance inside one
    storage (what interaction backward returns), give (base tensor, stride_table, stride_bag)."""
    g0 = grads[0]
    if g0.dim() != 2 or tuple(g0.shape) != (B, D) or g0.stride(1) != 1:
        return None
    st = g0.untyped_storage().data_ptr()
    sb = g0.stride(0)
    stt = (grads[1].storage_offset() - g0.storage_offset()) if len(grads) > 1 else B * D
    for k, g in enumerate(grads):
        if (g.untyped_storage().data_ptr() != st or g.dtype != g0.dtype or tuple(g.shape) != (B, D)
                or g.stride() != (sb, 1) or g.storage_offset() != g0.storage_offset() + k * stt):
            return None
    if stt <= 0 or sb <= 0:
        return None
    return g0, stt, sb


class _MergedEmbFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, trigger, module, indices, offsets, B):
        T, D = module.n_tables, module.dim
        s0 = 1 if module.reserve_dense_slot else 0
        pend, module._pending_fwd = module._pending_fwd, None
        if pend is not None and pend[0] is indices and pend[1] is offsets:
            buf = pend[2]  # pooled rows launched ahead on the side stream (prelaunch_forward)
            torch.cuda.current_stream().wait_event(pend[3])
            # that event is behind the previous step's update on the side stream: nothing left to join
            module._apply_event = None
            module._held_grad = None
        else:
            module.join()  # the previous step's update may still be running on the side stream
            buf = module._alloc_pooled(B, indices.device)
            with module._timed("emb_fwd"):
                ops.emb_fwd(module.tables, indices, offsets, B, buf[:, s0:], D, (T + s0) * D, module.error_flag)
            module._launch_prepare(indices, offsets, B)
        ctx.module, ctx.B = module, B
        ctx.save_for_backward(indices, offsets)
        ctx.set_materialize_grads(False)
        return tuple(buf[:, s0 + k] for k in range(T))

    @staticmethod
    def backward(ctx, *grads):
        module, B = ctx.module, ctx.B
        indices, offsets = ctx.saved_tensors
        T, D = module.n_tables, module.dim
        if all(g is None for g in grads):
            return (None,) * 5
        if any(g is None for g in grads):
            ref = next(g for g in grads if g is not None)
            grads = tuple(torch.zeros_like(ref) if g is None else g for g in grads)
        lay = _common_layout(grads, B, D)
        if lay is None:  # generic autograd graph: pack once (the fused DLRM path never gets here)
            packed = torch.stack([g.contiguous() for g in grads], dim=0)
            base, stt, sb = packed, B * D, D
        else:
            base, stt, sb = lay
        module._apply_update(indices, offsets, B, base, stt, sb)
        return (None,) * 5


class MergedEmbeddingBagWithSGD(nn.Module):
    """All tables of a DLRM in one module; ``forward`` pools, ``backward`` updates the rows."""

    def __init__(self, weights: List[torch.Tensor], lr: float = 0.01, weight_decay: float = 0.0,
                 optimizer: str = "sgd", eps: float = 1e-10):
        super().__init__()
        if weight_decay != 0.0:
            raise NotImplementedError("weight_decay != 0 is not supported by the fused update")
        dev = weights[0].device
        if dev.type != "cuda":
            raise RuntimeError("MergedEmbeddingBagWithSGD needs CUDA tensors (no CPU fallback)")
        self.n_tables = len(weights)
        self.dim = int(weights[0].shape[1])
        for k, w in enumerate(weights):
            self.add_module(str(k), _TableHolder(w))
        self.sgd_args = SGDArgs(bf16_trail=[None] * len(weights), weight_decay=0.0, lr=float(lr))
        self.optimizer_kind = optimizer
        self.eps = float(eps)
        self.output_dtype = None  # None: same dtype as the tables
        self.lr_tensor = None     # optional device scalar read at kernel time (CUDA-graph replay)
        self.error_flag = None    # optional int32[1] device flag set on an out-of-range index
        self.profile = None       # dict -> per-kernel CUDA-event pairs (see _timed)
        self.overlap_prepare = True
        # True: pooled buffer is [B, T+1, D] with slot 0 free for the dense feature (single-process
        # interaction reads it in place); False: [B, T, D] = the all-to-all send layout.
        self.reserve_dense_slot = True
        self._trigger = torch.zeros((), device=dev, requires_grad=True)
        self._ws = None
        self._side_stream = None
        self._prep_event = None
        # peer-memory (sharded) path: the NVLink-bound kernels run on the side stream next to the bottom MLP
        # (forward) and next to the bottom-MLP backward / gradient all-reduce / optimizer (update)
        self.overlap_sharded = True
        self._pending_sharded = {}    # chunk -> (indices, offsets, event) of a sharded forward launched ahead
        self._apply_event = None      # last fused update issued on the side stream
        self._chunk_ws = {}           # chunk -> backward workspace (micro-batched peer-memory path)
        # single-process path: the fused update runs on the side stream next to the bottom-MLP backward and the
        # dense optimizer step (both only need the interaction backward, like the update); join() orders readers
        self.overlap_update = True
        self._pending_fwd = None      # (indices, offsets, pooled buffer, event) of a forward launched ahead
        self._held_grad = None        # keeps the gradient buffer alive while the side stream reads it
        self._chunk_prep = {}         # chunk -> event of that chunk's index stage on the side stream
        self._rebuild_tables()
        if optimizer == "adagrad":
            self.tables.set_states([torch.zeros_like(w, dtype=torch.float32) for w in self.weights])
        elif optimizer != "sgd":
            raise ValueError(f"unknown optimizer {optimizer!r}")

    # ---- IPEX-shaped surface -------------------------------------------------------
    @classmethod
    def from_embeddingbag_list(cls, tables, lr: float = 0.01, weight_decay: float = 0.0, device=None,
                               optimizer: str = "sgd", eps: float = 1e-10):
        dev = torch.device(device) if device is not None else None
        weights = []
        for E in tables:
            if getattr(E, "mode", "sum") != "sum":
                raise NotImplementedError("only sum pooling is supported")
            w = E.weight.detach()
            if dev is None:
                dev = w.device if w.is_cuda else torch.device("cuda", torch.cuda.current_device())
            weights.append(w.to(dev).contiguous())
        return cls(weights, lr=lr, weight_decay=weight_decay, optimizer=optimizer, eps=eps)

    @property
    def weights(self) -> List[torch.Tensor]:
        self.join()  # readers on the current stream see the last fused update
        return self._raw_weights()

    def _raw_weights(self) -> List[torch.Tensor]:
        return [getattr(self, str(k)).weight.data for k in range(self.n_tables)]

    def state_dict(self, *args, **kwargs):
        self.join()
        return super().state_dict(*args, **kwargs)

    def optimizer_state_dict(self):
        """What the fused update keeps OUTSIDE the weights: Adagrad ``state_sum`` tables and the low-half trails of
        bf16 tables.  A checkpoint that drops them silently resets the accumulators / truncates the master weights."""
        self.join()
        return {"schema": 1, "optimizer": self.optimizer_kind, "lr": self.sgd_args.lr, "eps": self.eps,
                "states": None if self.tables.states is None else [s.clone() for s in self.tables.states],
                "trails": None if self.tables.trails is None else [t.clone() for t in self.tables.trails]}

    def load_optimizer_state_dict(self, sd):
        if sd.get("schema") != 1 or sd.get("optimizer") != self.optimizer_kind:
            raise RuntimeError("embedding optimizer state: wrong schema / optimizer kind")
        self.join()
        for name, cur in (("states", self.tables.states), ("trails", self.tables.trails)):
            saved = sd.get(name)
            if (saved is None) != (cur is None):
                raise RuntimeError(f"embedding optimizer state: {name} present on one side only")
            if saved is not None:
                for d, s_ in zip(cur, saved):
                    d.copy_(s_.to(d.device))

    def join(self):
        """Order the current stream after the last fused update issued on the side stream (no-op otherwise)."""
        ev = self._apply_event
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)
            self._apply_event = None
        self._held_grad = None

    def linearize_indices_and_offsets(self, indices: List[torch.Tensor], offsets: List[torch.Tensor],
                                      include_last_offsets: List[bool]):
        """Per-table (indices, offsets) -> (indices, offsets, indices_with_row_offsets):
        concatenated indices, one CSR array over all T*B bags with the trailing total, and
        global row ids."""
        dev = self._raw_weights()[0].device
        idx = [i.to(dev, torch.int64) for i in indices]
        offs = [o.to(dev, torch.int64) for o in offsets]
        # one launch for all tables (b200_linearize) instead of an add per table and three concatenations
        return ops.linearize(idx, offs, list(include_last_offsets), self.tables.d_row_base)

    def to_bfloat16_train(self, keep_trail: bool = True):
        """Store the tables in bf16: the forward gathers half the bytes.  With ``keep_trail`` (default,
        the reference's split-SGD storage, P/dlrm_s_pytorch.py:920-933) every table keeps the LOWER 16 bits
        of its fp32 master weights in a side tensor (``sgd_args.bf16_trail``): the fused update rebuilds the
        fp32 value, updates it exactly and splits it again, so no small update is ever rounded away.
        Without it the rows are plain bf16, rounded to nearest after each update."""
        trails = []
        for k in range(self.n_tables):
            h = getattr(self, str(k))
            w = h.weight.data
            if w.dtype != torch.float32:
                raise RuntimeError("to_bfloat16_train expects fp32 tables")
            if keep_trail:
                bits = w.view(torch.int32)
                trails.append((bits & 0xFFFF).to(torch.int16))             # low half (bit pattern)
                hi = (bits >> 16).to(torch.int16).view(torch.bfloat16)     # high half = truncated bf16
            else:
                hi = w.to(torch.bfloat16)
            h.weight = nn.Parameter(hi.contiguous(), requires_grad=False)
        self._rebuild_tables()
        if keep_trail:
            self.tables.set_trails(trails)
            self.sgd_args = self.sgd_args._replace(bf16_trail=trails)
        return self

    def master_weights(self):
        """fp32 view of the tables (exact when the trail is kept)."""
        out = []
        for k, w in enumerate(self.weights):
            if w.dtype == torch.bfloat16 and self.tables.trails is not None:
                hi = w.view(torch.int16).to(torch.int32) << 16
                lo = self.tables.trails[k].to(torch.int32) & 0xFFFF
                out.append((hi | lo).view(torch.float32))
            else:
                out.append(w.float())
        return out

    def forward(self, input, need_linearize_indices_and_offsets=None):
        need = True if need_linearize_indices_and_offsets is None else bool(need_linearize_indices_and_offsets)
        if need:
            idx_l, off_l = input[0], input[1]
            inc = input[2] if len(input) > 2 else [False] * len(idx_l)
            indices, offsets, _ = self.linearize_indices_and_offsets(idx_l, off_l, inc)
        else:
            indices, offsets = input[0], input[1]
        B = (offsets.numel() - 1) // self.n_tables
        if torch.is_grad_enabled() and self.training:
            return list(_MergedEmbFn.apply(self._trigger, self, indices, offsets, B))
        T, D = self.n_tables, self.dim
        s0 = 1 if self.reserve_dense_slot else 0
        self.join()
        buf = self._alloc_pooled(B, indices.device)
        ops.emb_fwd(self.tables, indices, offsets, B, buf[:, s0:], D, (T + s0) * D, self.error_flag)
        return [buf[:, s0 + k] for k in range(T)]

    def _alloc_pooled(self, B, device):
        s0 = 1 if self.reserve_dense_slot else 0
        buf = torch.empty(B, self.n_tables + s0, self.dim, dtype=self.output_dtype or self.tables.dtype, device=device)
        if s0:
            _register_pooled(buf)
        return buf

    def prelaunch_forward(self, indices, offsets):
        """Queue the pooled forward (and the index stage of the backward) of already linearised bags on the side
        stream NOW, so that they run next to whatever the caller does next on the current stream (the bottom MLP);
        the following ``forward`` with the same (indices, offsets) tensors picks the result up.  Training mode only;
        does nothing while per-kernel profiling is on."""
        if not (self.overlap_prepare and self.overlap_update and self.profile is None
                and torch.is_grad_enabled() and self.training):
            return
        B = (offsets.numel() - 1) // self.n_tables
        T, D = self.n_tables, self.dim
        s0 = 1 if self.reserve_dense_slot else 0
        main = torch.cuda.current_stream()
        if self._side_stream is None:
            self._side_stream = torch.cuda.Stream(device=indices.device)
        side = self._side_stream
        buf = self._alloc_pooled(B, indices.device)
        ws = self._workspace(indices.numel())
        side.wait_stream(main)  # inputs ready; the previous step's update is already ahead of us on `side`
        with torch.cuda.stream(side):
            ops.emb_fwd(self.tables, indices, offsets, B, buf[:, s0:], D, (T + s0) * D, self.error_flag)
            ev = torch.cuda.Event()
            ev.record(side)
            ops.emb_bwd_prepare(self.tables, indices, offsets, B, ws, self.error_flag)
            pe = torch.cuda.Event()
            pe.record(side)
        self._prep_event = pe
        self._pending_fwd = (indices, offsets, buf, ev)

    # ---- internals -----------------------------------------------------------------
    def _rebuild_tables(self):
        old = getattr(self, "tables", None)
        states = old.states if old is not None else None
        trails = old.trails if old is not None else None
        self.tables = ops.TableSet(self._raw_weights(), states)
        if trails is not None and self.tables.dtype == torch.bfloat16:
            self.tables.set_trails(trails)

    def _apply(self, fn, *a, **k):  # keep the device pointer table valid after .to()/.cuda()
        r = super()._apply(fn, *a, **k)
        if getattr(self, "tables", None) is not None:
            self._rebuild_tables()
        return r

    def _workspace(self, nnz: int, chunk: int = 0) -> torch.Tensor:
        if chunk != 0:  # every micro-batch in flight needs its own sort / dedup results
            ws = self._chunk_ws.get(chunk)
            need = ops.emb_bwd_workspace_bytes(nnz, self.dim)
            if ws is None or ws.numel() < need:
                if ws is not None and self._side_stream is not None:
                    ws.record_stream(self._side_stream)
                ws = self._chunk_ws[chunk] = torch.empty(need, dtype=torch.uint8, device=self._raw_weights()[0].device)
            return ws
        need = ops.emb_bwd_workspace_bytes(nnz, self.dim)
        if self._ws is None or self._ws.numel() < need:
            old = self._ws
            if old is not None and self._side_stream is not None:
                # a fused update of the previous step may still be reading the old block on the side stream: tell
                # the caching allocator, so that the block is not handed out again before that work has finished
                old.record_stream(self._side_stream)
            # grow geometrically (ragged bags: nnz changes from step to step)
            self._ws = torch.empty(max(need, 0 if old is None else old.numel() * 3 // 2), dtype=torch.uint8,
                                   device=self._raw_weights()[0].device)
        return self._ws

    @contextlib.contextmanager
    def _timed(self, name):
        """CUDA-event bracket on the launching stream, active only when ``self.profile`` is a dict
        (bench.py's roofline measurement); otherwise free."""
        if self.profile is None:
            yield
            return
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # keep the GPU busy for ~0.2 ms first: the host then runs ahead of the device, and the interval
        # a -> b holds the kernel(s) only, not the Python / ctypes time between the two record calls
        torch.cuda._sleep(400_000)
        a.record()
        yield
        b.record()
        self.profile.setdefault(name, []).append((a, b))

    def _set_prep(self, chunk, ev):
        if chunk == 0:
            self._prep_event = ev
        else:
            self._chunk_prep[chunk] = ev

    def _take_prep(self, chunk):
        if chunk == 0:
            ev, self._prep_event = self._prep_event, None
            return ev
        return self._chunk_prep.pop(chunk, None)

    def _take_pending_sharded(self, chunk):
        return self._pending_sharded.pop(chunk, None)

    def _launch_prepare(self, indices, offsets, B, chunk: int = 0):
        """Index stage of the backward (keys -> stable sort -> dedup).  It depends only on this
        step's indices, so it is issued on a side stream right after the forward gather and
        overlaps the interaction / MLP forward+backward; ``_apply_update`` joins it."""
        ws = self._workspace(indices.numel(), chunk)
        if not self.overlap_prepare:
            self._set_prep(chunk, None)
            return
        main = torch.cuda.current_stream()
        if self._side_stream is None:
            self._side_stream = torch.cuda.Stream(device=indices.device)
        self._side_stream.wait_stream(main)
        with torch.cuda.stream(self._side_stream):
            ops.emb_bwd_prepare(self.tables, indices, offsets, B, ws, self.error_flag)
            ev = torch.cuda.Event()
            ev.record(self._side_stream)
        self._set_prep(chunk, ev)

    def _launch_forward_sharded_async(self, chunks, feat0):
        """Peer-memory path: issue the pooled forward of every micro-batch (rows written straight into the
        destination ranks' T) and then their index stages on the side stream NOW -- the caller runs the bottom MLP
        next on the current stream, the sharded autograd function waits for the recorded event before its barrier.
        ``chunks``: list of (indices, offsets, PeerExchange)."""
        main = torch.cuda.current_stream()
        dev = chunks[0][0].device
        if self._side_stream is None:
            self._side_stream = torch.cuda.Stream(device=dev)
        side = self._side_stream
        wss = [self._workspace(idx.numel(), exch.chunk) for idx, _, exch in chunks]
        side.wait_stream(main)  # batch ready, last step's barriers passed; the previous update precedes on `side`
        with torch.cuda.stream(side):
            for idx, off, exch in chunks:  # every chunk reads the tables BEFORE any chunk's update (queued later)
                ops.emb_fwd_sharded(self.tables, idx, off, exch.lN * exch.world, exch.T_ptrs_dev, exch.lN, exch.F,
                                    feat0, exch.dtype, self.error_flag, first_dest=(exch.rank + 1) % exch.world)
                ev = torch.cuda.Event()
                ev.record(side)
                self._pending_sharded[exch.chunk] = (idx, off, ev)
            for (idx, off, exch), ws in zip(chunks, wss):
                ops.emb_bwd_prepare(self.tables, idx, off, exch.lN * exch.world, ws, self.error_flag)
                pe = torch.cuda.Event()
                pe.record(side)
                self._set_prep(exch.chunk, pe)
        self._apply_event = None  # the update before this forward is ordered by the side stream itself

    def _apply_update_sharded(self, indices, offsets, B, exch, feat0):
        """Backward of the peer-memory path (cloudtik_b200.p2p): gradients are read from every rank's dT."""
        opt = ops.OPT_ADAGRAD if self.optimizer_kind == "adagrad" else ops.OPT_SGD
        chunk = exch.chunk
        ws = self._workspace(indices.numel(), chunk)
        prep = self._take_prep(chunk)
        # multi-hot bags: stage the peers' gradient rows locally first (see PeerExchange.gather_grads)
        staged = indices.numel() >= 2 * self.n_tables * B

        def apply():
            if staged:
                G = exch.gather_grads(feat0, self.n_tables)
                ops.emb_bwd_apply(self.tables, indices.numel(), B, G, self.dim, self.n_tables * self.dim,
                                  self.sgd_args.lr, ws, optimizer=opt, eps=self.eps, lr_tensor=self.lr_tensor)
            else:
                ops.emb_bwd_apply_sharded(self.tables, indices.numel(), B, exch.dT_ptrs_dev, exch.lN, exch.F, feat0,
                                          exch.dtype, self.sgd_args.lr, ws, optimizer=opt, eps=self.eps,
                                          lr_tensor=self.lr_tensor)
        if self.overlap_sharded and prep is not None and self.profile is None:
            # NVLink-bound gather of the peers' gradient rows: run it on the side stream (already behind the index
            # stage) while the current stream does the bottom-MLP backward, the all-reduce and the optimizer -- or,
            # in a micro-batched step, the next chunk's forward / backward
            side = self._side_stream
            side.wait_stream(torch.cuda.current_stream())  # dT complete on every rank (barrier B2 is before us)
            with torch.cuda.stream(side):
                apply()
                ev = torch.cuda.Event()
                ev.record(side)
            self._apply_event = ev
            return
        if prep is not None:
            torch.cuda.current_stream().wait_event(prep)
        else:
            ops.emb_bwd_prepare(self.tables, indices, offsets, B, ws, self.error_flag)
        with self._timed("emb_bwd_apply"):
            apply()

    def _apply_update(self, indices, offsets, B, grad_base, stride_table, stride_bag):
        opt = ops.OPT_ADAGRAD if self.optimizer_kind == "adagrad" else ops.OPT_SGD
        ws = self._workspace(indices.numel())
        if self.overlap_update and self._prep_event is not None and self.profile is None:
            # HBM-bound update on the side stream (already behind the index stage there); the current stream goes
            # on with the bottom-MLP backward and the dense optimizer step, which need nothing from it
            side = self._side_stream
            side.wait_stream(torch.cuda.current_stream())  # the gradient rows are complete
            self._prep_event = None
            with torch.cuda.stream(side):
                ops.emb_bwd_apply(self.tables, indices.numel(), B, grad_base, stride_table, stride_bag,
                                  self.sgd_args.lr, ws, optimizer=opt, eps=self.eps, lr_tensor=self.lr_tensor)
                ev = torch.cuda.Event()
                ev.record(side)
            self._apply_event = ev
            self._held_grad = grad_base  # freed on the current stream only after join()
            return
        if self._prep_event is not None:
            torch.cuda.current_stream().wait_event(self._prep_event)
            self._prep_event = None
        else:
            with self._timed("emb_bwd_prepare"):
                ops.emb_bwd_prepare(self.tables, indices, offsets, B, ws, self.error_flag)
        with self._timed("emb_bwd_apply"):
            ops.emb_bwd_apply(self.tables, indices.numel(), B, grad_base, stride_table, stride_bag,
                              self.sgd_args.lr, ws, optimizer=opt, eps=self.eps, lr_tensor=self.lr_tensor)
