"""Small embedding tables replicated on every rank (data-parallel), next to the table-wise sharded large ones.

The reference's training variant splits the tables at ``--sparse-dense-boundary`` (default 2048 rows,
``T/dlrm_s_pytorch.py:266-269``): a table with fewer rows is built with ``sparse=False``, kept on EVERY rank
(``emb_dense``), looked up for the rank's LOCAL batch only (``T/dlrm_s_pytorch.py:489-509``) and wrapped in DDP so
that its dense gradient is all-reduced like the MLPs' (``T/dlrm_s_pytorch.py:988-997``); only the tables at or above
the boundary are sharded and exchanged.  On the Criteo-Terabyte shape that takes 10 of the 26 tables (3 ... 1543
rows) out of the pooled-embedding all-to-all: the exchange moves 16 tables instead of 26 and, at 8 ranks, every
rank owns exactly 2 of them.

Here the replicated tables are dense ``nn.Parameter``s ``{k}.weight`` (the optimizer that steps the MLPs steps them
too, the flat gradient all-reduce covers them), pooled by the same ``b200_emb_fwd`` kernel straight into the
interaction input, and their dense gradient comes from ``b200_emb_bwd_prepare`` + ``b200_emb_bwd_dense_grad``
(windowed segment sums: a 3-row table receives a third of the batch per row).
"""
from __future__ import annotations

from typing import List

import torch
import torch.nn as nn

from . import ops


class _Table(nn.Module):
    def __init__(self, weight: torch.Tensor):
        super().__init__()
        self.weight = nn.Parameter(weight, requires_grad=True)


class ReplicatedEmbeddingBags(nn.Module):
    def __init__(self, weights: List[torch.Tensor]):
        super().__init__()
        if not weights:
            raise ValueError("ReplicatedEmbeddingBags needs at least one table")
        for k, w in enumerate(weights):
            if not w.is_cuda or w.dtype != torch.float32:
                raise RuntimeError("replicated tables are fp32 CUDA tensors (no CPU path)")
            self.add_module(str(k), _Table(w.contiguous()))
        self.n_tables = len(weights)
        self.dim = int(weights[0].shape[1])
        self.error_flag = None
        self._ws = None
        self._prep_event = None
        # The index stage and (when `defer` is set by the step driver) the dense-gradient kernels run on this
        # module's own low-priority stream: nothing on the step's main chain needs them before the gradient
        # all-reduce, which `join()` precedes.
        self._stream = None
        self._grad_event = None
        self._pending_grads = None
        self._pool_event = None
        self.defer = False
        self._rebuild()

    def params(self) -> List[nn.Parameter]:
        return [getattr(self, str(k)).weight for k in range(self.n_tables)]

    def _rebuild(self):
        self.tables = ops.TableSet([p.data for p in self.params()])

    def _apply(self, fn, *a, **k):
        r = super()._apply(fn, *a, **k)
        if getattr(self, "tables", None) is not None:
            self._rebuild()
        return r

    def _tables_checked(self):
        # the optimizer updates the parameters in place, so the pointer table stays valid; a re-assigned
        # .data (load_state_dict assigns in place too, but be safe) is picked up here
        if any(p.data_ptr() != w.data_ptr() for p, w in zip(self.params(), self.tables.weights)):
            self._rebuild()
        return self.tables

    def _workspace(self, nnz: int):
        need = ops.emb_bwd_workspace_bytes(nnz, self.dim)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.tables.device)
        return self._ws

    def pool_into(self, indices, offsets, B: int, out: torch.Tensor, stride_table: int, stride_bag: int):
        """Pooled rows of the B local bags of every replicated table into ``out`` (a view of the interaction
        input: table t of bag b at ``out.data_ptr + (t * stride_table + b * stride_bag)`` elements)."""
        ops.emb_fwd(self._tables_checked(), indices, offsets, B, out, stride_table, stride_bag, self.error_flag)

    def pool_into_async(self, indices, offsets, B: int, out: torch.Tensor, stride_table: int, stride_bag: int,
                        prepare: bool = True):
        """``pool_into`` (+ the index stage of the backward) on this module's own stream, to be called before the
        caller runs something else on the current stream (the bottom MLP); ``wait_pool()`` orders the current stream
        behind the lookup.  ``out`` must not be read by anything still queued on the current stream."""
        side = self._own_stream()
        side.wait_stream(torch.cuda.current_stream())
        tables = self._tables_checked()
        with torch.cuda.stream(side):
            ops.emb_fwd(tables, indices, offsets, B, out, stride_table, stride_bag, self.error_flag)
            ev = torch.cuda.Event()
            ev.record(side)
            self._pool_event = (indices, offsets, ev)
            if prepare:
                ops.emb_bwd_prepare(tables, indices, offsets, B, self._workspace(indices.numel()), self.error_flag)
                pe = torch.cuda.Event()
                pe.record(side)
                self._prep_event = pe

    def wait_pool(self, indices, offsets) -> bool:
        """True (and the current stream now waits for it) if ``pool_into_async`` ran for these very tensors."""
        pend, self._pool_event = self._pool_event, None
        if pend is not None and pend[0] is indices and pend[1] is offsets:
            torch.cuda.current_stream().wait_event(pend[2])
            return True
        return False

    def _own_stream(self):
        if self._stream is None:
            # high priority: the lookup feeds the interaction, the dense gradients feed the end-of-step all-reduce --
            # queued behind the thousands of blocks of the (low-priority) NVLink-bound update they arrive 200 us late
            self._stream = torch.cuda.Stream(device=self.tables.device, priority=-1)
        return self._stream

    def launch_prepare(self, indices, offsets, B: int):
        """Index stage of the backward (keys, stable sort, dedup) on this module's own stream: it only needs the
        indices, so it runs next to the forward pass."""
        ws = self._workspace(indices.numel())
        tables = self._tables_checked()
        side = self._own_stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            ops.emb_bwd_prepare(tables, indices, offsets, B, ws, self.error_flag)
            ev = torch.cuda.Event()
            ev.record(side)
        self._prep_event = ev

    def dense_grads(self, indices, offsets, B: int, grad_base: torch.Tensor, stride_table: int, stride_bag: int):
        """Per-table dense fp32 gradients (row-slices of one [total_rows, D] buffer) from the gradient of the
        pooled rows, addressed like ``pool_into``'s output.  ``launch_prepare`` must have run for these bags.
        With ``defer`` set the kernels are queued on this module's stream and ``join()`` must be called on the
        consuming stream before the gradients are read (cloudtik_b200.dlrm.train_step_chunked does)."""
        tables = self.tables
        ws = self._workspace(indices.numel())

        def run():
            G = torch.zeros(tables.total_rows, self.dim, dtype=torch.float32, device=tables.device)
            ops.emb_bwd_dense_grad(tables, indices.numel(), B, grad_base, stride_table, stride_bag, G, ws)
            return G
        if self.defer:
            side = self._own_stream()  # already behind the index stage
            side.wait_stream(torch.cuda.current_stream())  # the gradient of the pooled rows is complete
            self._prep_event = None
            with torch.cuda.stream(side):
                G = run()
                ev = torch.cuda.Event()
                ev.record(side)
            self._grad_event = ev
            # autograd must not touch these tensors before join(): hand them to the parameters there, not now
            views, base = [], 0
            for n in tables.rows:
                views.append(G[base:base + n])
                base += n
            self._pending_grads = views
            return [None] * len(views)
        else:
            if self._prep_event is not None:
                torch.cuda.current_stream().wait_event(self._prep_event)
                self._prep_event = None
            G = run()
        out, base = [], 0
        for n in tables.rows:
            out.append(G[base:base + n])
            base += n
        return out

    def join(self):
        """Order the current stream behind deferred dense-gradient kernels (no-op otherwise)."""
        ev, self._grad_event = self._grad_event, None
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)
            for p, g in zip(self.params(), self._pending_grads):
                p.grad = g if p.grad is None else p.grad + g
            self._pending_grads = None
