"""Functional wrappers: torch CUDA tensors -> C-ABI calls (pointers, sizes, current stream).

torch is used here only for device memory and streams.  Every function raises
``B200Error`` on a non-zero status; nothing falls back to torch or to the CPU.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import BF16, F32, OPT_ADAGRAD, OPT_SGD, check, lib


def _dt(t: torch.dtype) -> int:
    if t == torch.float32:
        return F32
    if t == torch.bfloat16:
        return BF16
    raise _lib.B200Error(f"unsupported dtype {t} (float32 / bfloat16 only)")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _cuda(t: torch.Tensor, name: str, dtype=None):
    if not t.is_cuda:
        raise _lib.B200Error(f"{name} must be a CUDA tensor (no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise _lib.B200Error(f"{name} must be {dtype}, got {t.dtype}")
    return t


class TableSet:
    """Device-side description of a list of embedding tables [rows_t, D] (same D, same dtype)."""

    def __init__(self, weights, states=None):
        assert len(weights) > 0
        w0 = weights[0]
        self.weights = list(weights)
        self.D = int(w0.shape[1])
        self.dtype = w0.dtype
        self.device = w0.device
        for w in weights:
            _cuda(w, "weight")
            if w.dtype != self.dtype or w.shape[1] != self.D or not w.is_contiguous():
                raise _lib.B200Error("all tables must be contiguous with the same dim and dtype")
            if w.data_ptr() % 16:
                raise _lib.B200Error("table base pointers must be 16-byte aligned")
        rows = [int(w.shape[0]) for w in weights]
        self.rows = rows
        self.total_rows = sum(rows)
        base = [0]
        for r in rows:
            base.append(base[-1] + r)
        self.d_ptrs = torch.tensor([w.data_ptr() for w in weights], dtype=torch.int64, device=self.device)
        self.d_rows = torch.tensor(rows, dtype=torch.int64, device=self.device)
        self.d_row_base = torch.tensor(base, dtype=torch.int64, device=self.device)
        self.states = None
        self.d_state_ptrs = None
        self.trails = None
        self.d_trail_ptrs = None
        if states is not None:
            self.set_states(states)

    def set_trails(self, trails):
        """bf16 tables: int16/uint16-bit-pattern [rows, D] tensors holding the low halves of the fp32 masters."""
        for t_, w in zip(trails, self.weights):
            assert t_.is_cuda and t_.shape == w.shape and t_.is_contiguous() and t_.element_size() == 2
        self.trails = list(trails)
        self.d_trail_ptrs = torch.tensor([t_.data_ptr() for t_ in trails], dtype=torch.int64, device=self.device)

    def set_states(self, states):
        for s, w in zip(states, self.weights):
            _cuda(s, "state", torch.float32)
            assert s.shape == w.shape and s.is_contiguous()
        self.states = list(states)
        self.d_state_ptrs = torch.tensor([s.data_ptr() for s in states], dtype=torch.int64, device=self.device)

    def __len__(self):
        return len(self.weights)


def emb_fwd(tables: TableSet, indices, offsets, B: int, out: torch.Tensor, stride_table: int,
            stride_bag: int, error: torch.Tensor | None = None):
    """out[t*stride_table + b*stride_bag + :D] = sum of rows of bag (t, b).  See b200_emb_fwd."""
    _cuda(indices, "indices", torch.int64)
    _cuda(offsets, "offsets", torch.int64)
    _cuda(out, "out")
    T = len(tables)
    if offsets.numel() != T * B + 1:
        raise _lib.B200Error(f"offsets must have T*B+1 = {T * B + 1} entries, got {offsets.numel()}")
    check(lib.b200_emb_fwd(tables.d_ptrs.data_ptr(), tables.d_rows.data_ptr(), indices.data_ptr(),
                           offsets.data_ptr(), out.data_ptr(), indices.numel(), T, B, tables.D,
                           stride_table, stride_bag, _dt(tables.dtype), _dt(out.dtype),
                           error.data_ptr() if error is not None else None, _stream()), "b200_emb_fwd")
    return out


def sort_workspace_bytes(n: int) -> int:
    return int(lib.b200_sort_workspace_bytes(n))


def sort_dedup(keys: torch.Tensor, vals: torch.Tensor, key_bound: int):
    """Stable sort of (key, val) uint32 pairs (given as int32/uint32-bit-pattern tensors) and
    run-length dedup.  Returns (sorted_keys, sorted_vals, unique, seg_offsets, num_unique) with
    unique / seg_offsets trimmed on the host (this helper synchronises; the fused backward does not)."""
    n = keys.numel()
    dev = keys.device
    _cuda(keys, "keys", torch.int32)
    _cuda(vals, "vals", torch.int32)
    skeys = torch.empty(n, dtype=torch.int32, device=dev)
    svals = torch.empty(n, dtype=torch.int32, device=dev)
    uniq = torch.empty(n, dtype=torch.int32, device=dev)
    seg = torch.empty(n + 1, dtype=torch.int32, device=dev)
    nu = torch.zeros(1, dtype=torch.int32, device=dev)
    ws_bytes = sort_workspace_bytes(n)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    check(lib.b200_sort_dedup(keys.data_ptr(), vals.data_ptr(), n, key_bound, skeys.data_ptr(),
                              svals.data_ptr(), uniq.data_ptr(), seg.data_ptr(), nu.data_ptr(),
                              ws.data_ptr(), ws_bytes, _stream()), "b200_sort_dedup")
    U = int(nu.item())
    return skeys, svals, uniq[:U], seg[:U + 1], U


def emb_bwd_workspace_bytes(nnz: int, D: int) -> int:
    return int(lib.b200_emb_bwd_workspace_bytes(nnz, D))


def emb_bwd_update(tables: TableSet, indices, offsets, B: int, grad: torch.Tensor, stride_table: int,
                   stride_bag: int, lr: float, ws: torch.Tensor, optimizer: int = OPT_SGD,
                   eps: float = 1e-10, lr_tensor: torch.Tensor | None = None,
                   error: torch.Tensor | None = None):
    """Fused sparse backward + row update.  See b200_emb_bwd_update."""
    _cuda(indices, "indices", torch.int64)
    _cuda(offsets, "offsets", torch.int64)
    _cuda(grad, "grad")
    T = len(tables)
    if optimizer == OPT_ADAGRAD and tables.d_state_ptrs is None:
        raise _lib.B200Error("Adagrad needs per-table fp32 state (TableSet.set_states)")
    check(lib.b200_emb_bwd_update(
        tables.d_ptrs.data_ptr(),
        tables.d_state_ptrs.data_ptr() if tables.d_state_ptrs is not None else None,
        tables.d_trail_ptrs.data_ptr() if tables.d_trail_ptrs is not None else None,
        tables.d_rows.data_ptr(), tables.d_row_base.data_ptr(), indices.data_ptr(), offsets.data_ptr(),
        indices.numel(), grad.data_ptr(), T, B, tables.D, stride_table, stride_bag, _dt(tables.dtype),
        _dt(grad.dtype), optimizer, float(lr), lr_tensor.data_ptr() if lr_tensor is not None else None,
        float(eps), tables.total_rows, ws.data_ptr(), ws.numel() * ws.element_size(),
        error.data_ptr() if error is not None else None, _stream()), "b200_emb_bwd_update")


def emb_bwd_prepare(tables: TableSet, indices, offsets, B: int, ws: torch.Tensor,
                    error: torch.Tensor | None = None, nnz_valid: torch.Tensor | None = None):
    """Index stage of the backward (keys, stable sort, dedup) into ``ws``.  See b200_emb_bwd_prepare.
    ``nnz_valid``: device int64 scalar -- ``indices`` is a fixed-capacity array filled up to that count
    (b200_emb_bwd_prepare_partial)."""
    _cuda(indices, "indices", torch.int64)
    _cuda(offsets, "offsets", torch.int64)
    if nnz_valid is not None:
        _cuda(nnz_valid, "nnz_valid", torch.int64)
        check(lib.b200_emb_bwd_prepare_partial(
            tables.d_rows.data_ptr(), tables.d_row_base.data_ptr(), indices.data_ptr(), offsets.data_ptr(),
            indices.numel(), nnz_valid.data_ptr(), len(tables), B, tables.D, tables.total_rows, ws.data_ptr(),
            ws.numel() * ws.element_size(), error.data_ptr() if error is not None else None, _stream()),
            "b200_emb_bwd_prepare_partial")
        return
    check(lib.b200_emb_bwd_prepare(tables.d_rows.data_ptr(), tables.d_row_base.data_ptr(), indices.data_ptr(),
                                   offsets.data_ptr(), indices.numel(), len(tables), B, tables.D,
                                   tables.total_rows, ws.data_ptr(), ws.numel() * ws.element_size(),
                                   error.data_ptr() if error is not None else None, _stream()),
          "b200_emb_bwd_prepare")


def rowwise_filter(indices, offsets, world: int, rank: int, capacity: int, error: torch.Tensor | None = None,
                   out=None):
    """Bags over global row ids -> bags over the rows rank ``rank`` of ``world`` owns (id % world == rank, local id =
    id // world) in a FIXED-capacity index array; ``offsets[-1]`` of the result is the fill level (it stays on the
    device).  Returns (local_indices int64[capacity], local_offsets int64[n_bags + 1]).  See b200_rowwise_filter."""
    _cuda(indices, "indices", torch.int64)
    _cuda(offsets, "offsets", torch.int64)
    nb = offsets.numel() - 1
    dev = indices.device
    if out is None:
        out = (torch.empty(max(capacity, 1), dtype=torch.int64, device=dev),
               torch.empty(nb + 1, dtype=torch.int64, device=dev))
    li, lo = out
    nbytes = int(lib.b200_rowwise_filter_workspace_bytes(nb))
    ws = torch.empty(max(nbytes, 8), dtype=torch.uint8, device=dev)
    check(lib.b200_rowwise_filter(indices.data_ptr(), offsets.data_ptr(), nb, world, rank, li.data_ptr(), capacity,
                                  lo.data_ptr(), error.data_ptr() if error is not None else None, ws.data_ptr(),
                                  nbytes, _stream()), "b200_rowwise_filter")
    return li, lo


def emb_bwd_apply(tables: TableSet, nnz: int, B: int, grad: torch.Tensor, stride_table: int, stride_bag: int,
                  lr: float, ws: torch.Tensor, optimizer: int = OPT_SGD, eps: float = 1e-10,
                  lr_tensor: torch.Tensor | None = None):
    """Segment-reduce + row update from a prepared ``ws``.  See b200_emb_bwd_apply."""
    _cuda(grad, "grad")
    if optimizer == OPT_ADAGRAD and tables.d_state_ptrs is None:
        raise _lib.B200Error("Adagrad needs per-table fp32 state (TableSet.set_states)")
    check(lib.b200_emb_bwd_apply(
        tables.d_ptrs.data_ptr(), tables.d_state_ptrs.data_ptr() if tables.d_state_ptrs is not None else None,
        tables.d_trail_ptrs.data_ptr() if tables.d_trail_ptrs is not None else None,
        tables.d_row_base.data_ptr(), nnz, grad.data_ptr(), len(tables), B, tables.D, stride_table, stride_bag,
        _dt(tables.dtype), _dt(grad.dtype), optimizer, float(lr),
        lr_tensor.data_ptr() if lr_tensor is not None else None, float(eps), ws.data_ptr(),
        ws.numel() * ws.element_size(), _stream()), "b200_emb_bwd_apply")


def emb_fwd_sharded(tables: TableSet, indices, offsets, B: int, out_ptrs_dev: int, lN: int, F: int, feat0: int,
                    out_dtype: torch.dtype, error: torch.Tensor | None = None, accumulate: bool = False,
                    first_dest: int = 0):
    """Pooled lookup of this rank's tables for the GLOBAL batch, written straight into every rank's
    interaction input over peer memory.  ``out_ptrs_dev``: device address of the [world] pointer table.
    ``accumulate``: ADD into the (zeroed) destination instead of storing -- row-wise sharded tables.
    See b200_emb_fwd_sharded / b200_emb_fwd_sharded_add."""
    _cuda(indices, "indices", torch.int64)
    _cuda(offsets, "offsets", torch.int64)
    # first_dest: rank whose bags are visited first ((own rank + 1) % world spreads the ranks over the destinations)
    check(lib.b200_emb_fwd_sharded_from(tables.d_ptrs.data_ptr(), tables.d_rows.data_ptr(), indices.data_ptr(),
                                        offsets.data_ptr(), out_ptrs_dev, indices.numel(), len(tables), B, tables.D,
                                        lN, F, feat0, _dt(tables.dtype), _dt(out_dtype), 1 if accumulate else 0,
                                        int(first_dest), error.data_ptr() if error is not None else None, _stream()),
          "b200_emb_fwd_sharded")


def emb_bwd_apply_sharded(tables: TableSet, nnz: int, B: int, grad_ptrs_dev: int, lN: int, F: int, feat0: int,
                          grad_dtype: torch.dtype, lr: float, ws: torch.Tensor, optimizer: int = OPT_SGD,
                          eps: float = 1e-10, lr_tensor: torch.Tensor | None = None):
    """Row update from a prepared ``ws`` with the gradient rows read from every rank's dT over peer memory.
    See b200_emb_bwd_apply_sharded."""
    if optimizer == OPT_ADAGRAD and tables.d_state_ptrs is None:
        raise _lib.B200Error("Adagrad needs per-table fp32 state (TableSet.set_states)")
    check(lib.b200_emb_bwd_apply_sharded(
        tables.d_ptrs.data_ptr(), tables.d_state_ptrs.data_ptr() if tables.d_state_ptrs is not None else None,
        tables.d_trail_ptrs.data_ptr() if tables.d_trail_ptrs is not None else None,
        tables.d_row_base.data_ptr(), nnz, grad_ptrs_dev, len(tables), B, tables.D, lN, F, feat0,
        _dt(tables.dtype), _dt(grad_dtype), optimizer, float(lr),
        lr_tensor.data_ptr() if lr_tensor is not None else None, float(eps), ws.data_ptr(),
        ws.numel() * ws.element_size(), _stream()), "b200_emb_bwd_apply_sharded")


def emb_bwd_coalesce(tables: TableSet, indices, offsets, B: int, grad: torch.Tensor, stride_table: int,
                     stride_bag: int, ws: torch.Tensor, error: torch.Tensor | None = None):
    """Coalesced sparse gradient of the pooled lookup: returns (unique_rows int64[U] global row ids ascending,
    values fp32[U, D]).  Synchronises once to read U (a sparse tensor needs a concrete nnz).
    See b200_emb_bwd_coalesce."""
    _cuda(indices, "indices", torch.int64)
    _cuda(offsets, "offsets", torch.int64)
    _cuda(grad, "grad")
    nnz = indices.numel()
    dev = indices.device
    rows = torch.empty(nnz, dtype=torch.int64, device=dev)
    vals = torch.empty(nnz, tables.D, dtype=torch.float32, device=dev)
    nu = torch.zeros(1, dtype=torch.int32, device=dev)
    check(lib.b200_emb_bwd_coalesce(tables.d_rows.data_ptr(), tables.d_row_base.data_ptr(), indices.data_ptr(),
                                    offsets.data_ptr(), nnz, grad.data_ptr(), len(tables), B, tables.D,
                                    stride_table, stride_bag, _dt(grad.dtype), tables.total_rows, rows.data_ptr(),
                                    vals.data_ptr(), nu.data_ptr(), ws.data_ptr(), ws.numel() * ws.element_size(),
                                    error.data_ptr() if error is not None else None, _stream()),
          "b200_emb_bwd_coalesce")
    U = int(nu.item())
    return rows[:U], vals[:U]


def emb_bwd_dense_grad(tables: TableSet, nnz: int, B: int, grad: torch.Tensor, stride_table: int, stride_bag: int,
                       out: torch.Tensor, ws: torch.Tensor):
    """Dense fp32 gradient ``out[total_rows, D]`` of the pooled lookup from a prepared ``ws`` (rows no bag touched
    keep what ``out`` held: zero it first).  See b200_emb_bwd_dense_grad."""
    _cuda(grad, "grad")
    _cuda(out, "out", torch.float32)
    if tuple(out.shape) != (tables.total_rows, tables.D) or not out.is_contiguous():
        raise _lib.B200Error("emb_bwd_dense_grad: out must be a contiguous [total_rows, D] fp32 tensor")
    check(lib.b200_emb_bwd_dense_grad(tables.d_row_base.data_ptr(), nnz, grad.data_ptr(), len(tables), B, tables.D,
                                      stride_table, stride_bag, _dt(grad.dtype), out.data_ptr(), ws.data_ptr(),
                                      ws.numel() * ws.element_size(), _stream()), "b200_emb_bwd_dense_grad")
    return out


def interact_fwd(T: torch.Tensor, out: torch.Tensor | None = None, pad_to: int = 1, itself: bool = False) -> torch.Tensor:
    """T [B, F, D] contiguous -> R [B, ldr] with ldr = D + F(F-1)/2 (``itself``: D + F(F+1)/2, the diagonal
    included) rounded up to ``pad_to`` (padding columns are written as zeros).  See b200_interact_fwd[_ex]."""
    _cuda(T, "T")
    if T.dim() != 3 or not T.is_contiguous():
        raise _lib.B200Error("interact_fwd: T must be a contiguous [B, F, D] tensor")
    B, F, D = T.shape
    W = D + F * (F - 1) // 2 + (F if itself else 0)
    if out is None:
        out = torch.empty(B, (W + pad_to - 1) // pad_to * pad_to, dtype=T.dtype, device=T.device)
    if out.dim() != 2 or out.shape[0] != B or out.shape[1] < W or out.stride(1) != 1 or out.dtype != T.dtype:
        raise _lib.B200Error("interact_fwd: bad output tensor")
    check(lib.b200_interact_fwd_ex(T.data_ptr(), out.data_ptr(), B, F, D, out.stride(0), _dt(T.dtype),
                                   1 if itself else 0, _stream()), "b200_interact_fwd")
    return out


def interact_bwd(T: torch.Tensor, dR: torch.Tensor, out: torch.Tensor | None = None, itself: bool = False) -> torch.Tensor:
    """dT [B, F, D] from T and dR [B, >= D + number of pairs] (row stride = dR.stride(0)).  See b200_interact_bwd[_ex]."""
    _cuda(T, "T")
    _cuda(dR, "dR", T.dtype)
    B, F, D = T.shape
    if (not T.is_contiguous() or dR.dim() != 2 or dR.shape[0] != B
            or dR.shape[1] < D + F * (F - 1) // 2 + (F if itself else 0) or dR.stride(1) != 1):
        raise _lib.B200Error("interact_bwd: bad shapes / non-contiguous input")
    if out is None:
        out = torch.empty_like(T)
    check(lib.b200_interact_bwd_ex(T.data_ptr(), dR.data_ptr(), out.data_ptr(), B, F, D, dR.stride(0), _dt(T.dtype),
                                   1 if itself else 0, _stream()), "b200_interact_bwd")
    return out


def linearize(idx_list, off_list, include_last, row_base: torch.Tensor | None, want_iwro: bool = True):
    """Per-table bags (device int64 tensors) -> (indices, offsets[, global row ids]) in the merged layout with ONE
    launch.  ``include_last[t]``: table t's offsets carry a trailing end offset (dropped).  See b200_linearize."""
    T = len(idx_list)
    dev = idx_list[0].device
    idx = [_cuda(i, "indices", torch.int64).contiguous() for i in idx_list]
    offs = []
    for o, inc in zip(off_list, include_last):
        o = _cuda(o, "offsets", torch.int64)
        offs.append((o[:-1] if inc else o).contiguous())
    B = offs[0].numel()
    if any(o.numel() != B for o in offs):
        raise _lib.B200Error("all tables must be indexed with the same number of bags")
    prefix = [0]
    for i in idx:
        prefix.append(prefix[-1] + i.numel())
    meta = torch.tensor([i.data_ptr() for i in idx] + [o.data_ptr() for o in offs] + prefix, dtype=torch.int64).to(
        dev, non_blocking=True)
    indices = torch.empty(prefix[-1], dtype=torch.int64, device=dev)
    offsets = torch.empty(T * B + 1, dtype=torch.int64, device=dev)
    iwro = torch.empty(prefix[-1], dtype=torch.int64, device=dev) if (want_iwro and row_base is not None) else None
    p0 = meta.data_ptr()
    check(lib.b200_linearize(p0, p0 + 8 * T, p0 + 16 * T, row_base.data_ptr() if iwro is not None else None, T, B,
                             max(i.numel() for i in idx), indices.data_ptr(), offsets.data_ptr(),
                             iwro.data_ptr() if iwro is not None else None, _stream()), "b200_linearize")
    # the per-table inputs and the pointer table must outlive the launch on this stream: torch's allocator is
    # stream-ordered, so tensors freed after the launch are not reused before it ran
    return indices, offsets, iwro


def criteo_transform(records: torch.Tensor, max_ind_range: int = 0, n_dense: int = 13, tab_begin: int = 0,
                     tab_count: int | None = None, row_begin: int = 0, row_count: int | None = None,
                     dense_stride: int | None = None, out=None):
    """Raw Criteo int32 records [N, tot_fea] (on the device) -> (dense fp32 [rows, dense_stride],
    label fp32 [rows, 1], indices int64 [tab_count * N], offsets int64 [tab_count * N + 1]).
    See b200_criteo_transform; ``out`` = the same 4-tuple preallocated (CUDA-graph replay)."""
    _cuda(records, "records", torch.int32)
    if records.dim() != 2 or not records.is_contiguous():
        raise _lib.B200Error("criteo_transform: records must be a contiguous [N, tot_fea] int32 tensor")
    N, tot = records.shape
    n_tab_all = tot - 1 - n_dense
    tab_count = n_tab_all - tab_begin if tab_count is None else tab_count
    row_count = N - row_begin if row_count is None else row_count
    dense_stride = n_dense if dense_stride is None else dense_stride
    if out is None:
        dev = records.device
        out = (torch.empty(row_count, dense_stride, dtype=torch.float32, device=dev),
               torch.empty(row_count, 1, dtype=torch.float32, device=dev),
               torch.empty(tab_count * N, dtype=torch.int64, device=dev),
               torch.empty(tab_count * N + 1, dtype=torch.int64, device=dev))
    dense, label, indices, offsets = out
    if (tuple(dense.shape) != (row_count, dense_stride) or label.numel() != row_count
            or indices.numel() != tab_count * N or offsets.numel() != tab_count * N + 1):
        raise _lib.B200Error("criteo_transform: preallocated outputs have the wrong shape")
    for t, dt in ((dense, torch.float32), (label, torch.float32), (indices, torch.int64), (offsets, torch.int64)):
        _cuda(t, "out", dt)
        if not t.is_contiguous():
            raise _lib.B200Error("criteo_transform: outputs must be contiguous")
    check(lib.b200_criteo_transform(records.data_ptr(), N, tot, n_dense, int(max_ind_range), tab_begin, tab_count,
                                    row_begin, row_count, dense.data_ptr(), dense_stride, label.data_ptr(),
                                    indices.data_ptr(), offsets.data_ptr(), _stream()), "b200_criteo_transform")
    return out


def relu_bwd_bias(dy: torch.Tensor, y: torch.Tensor):
    """(dz, db): dz = dy * (y > 0) in bf16, db = fp32 column sums of dz.  See b200_relu_bwd_bias."""
    _cuda(dy, "dy", torch.bfloat16)
    _cuda(y, "y", torch.bfloat16)
    if dy.dim() != 2 or dy.shape != y.shape or not dy.is_contiguous() or not y.is_contiguous():
        raise _lib.B200Error("relu_bwd_bias: dy and y must be contiguous [B, N] tensors of the same shape")
    B, N = dy.shape
    dz = torch.empty_like(dy)
    db = torch.empty(N, dtype=torch.float32, device=dy.device)
    nbytes = lib.b200_relu_bwd_bias_workspace_bytes(B, N)
    ws = torch.empty(max(nbytes, 4) // 4, dtype=torch.float32, device=dy.device)
    check(lib.b200_relu_bwd_bias(dy.data_ptr(), y.data_ptr(), dz.data_ptr(), db.data_ptr(), B, N, BF16,
                                 ws.data_ptr(), nbytes, _stream()), "b200_relu_bwd_bias")
    return dz, db


class QuantTableSet:
    """int8 tables [rows_t, D] with one fp32 scale per row (zero point 0): inference only."""

    def __init__(self, qweights, scales):
        assert len(qweights) > 0 and len(qweights) == len(scales)
        self.weights, self.scales = list(qweights), list(scales)
        self.D = int(qweights[0].shape[1])
        self.device = qweights[0].device
        for q, s in zip(qweights, scales):
            _cuda(q, "qweight", torch.int8)
            _cuda(s, "scale", torch.float32)
            if q.shape[1] != self.D or not q.is_contiguous() or s.numel() != q.shape[0] or not s.is_contiguous():
                raise _lib.B200Error("int8 tables: [rows, D] contiguous weights and one scale per row")
            if q.data_ptr() % 16:
                raise _lib.B200Error("table base pointers must be 16-byte aligned")
        self.rows = [int(q.shape[0]) for q in qweights]
        self.d_ptrs = torch.tensor([q.data_ptr() for q in qweights], dtype=torch.int64, device=self.device)
        self.d_scale_ptrs = torch.tensor([s.data_ptr() for s in scales], dtype=torch.int64, device=self.device)
        self.d_rows = torch.tensor(self.rows, dtype=torch.int64, device=self.device)

    def __len__(self):
        return len(self.weights)


def emb_fwd_q8(tables: QuantTableSet, indices, offsets, B: int, out: torch.Tensor, stride_table: int,
               stride_bag: int, error: torch.Tensor | None = None):
    """Pooled forward over int8 tables, dequantised on the fly.  See b200_emb_fwd_q8."""
    _cuda(indices, "indices", torch.int64)
    _cuda(offsets, "offsets", torch.int64)
    _cuda(out, "out")
    T = len(tables)
    if offsets.numel() != T * B + 1:
        raise _lib.B200Error(f"offsets must have T*B+1 = {T * B + 1} entries, got {offsets.numel()}")
    check(lib.b200_emb_fwd_q8(tables.d_ptrs.data_ptr(), tables.d_scale_ptrs.data_ptr(), tables.d_rows.data_ptr(),
                              indices.data_ptr(), offsets.data_ptr(), out.data_ptr(), indices.numel(), T, B,
                              tables.D, stride_table, stride_bag, _dt(out.dtype),
                              error.data_ptr() if error is not None else None, _stream()), "b200_emb_fwd_q8")
    return out
