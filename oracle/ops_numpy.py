"""numpy restatement of every operator on the DLRM training hot path.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``): used by ``tests/``, by
``__graft_entry__.smoke()`` and by ``bench.py``'s CPU-baseline leg as the checker.
Never imported by ``cloudtik_b200``.

Path abbreviations: ``P/`` = reference
``applications/ai/quickstart/models/recommendation/pytorch/dlrm/product``,
``T/`` = ``.../dlrm/training/bfloat16``.

The arithmetic of the reference lives in PyTorch ATen (torch==1.13.1 pinned by
``python/cloudtik/runtime/ai/scripts/install.sh:79``); the functions below restate
the operator semantics the reference's call sites rely on.  They are pinned in
``tests/test_oracle_golden.py`` against fixtures produced by the reference itself.
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------
# index bookkeeping (bit-exact stages)
# --------------------------------------------------------------------------


def bag_lengths(offsets: np.ndarray, nnz: int) -> np.ndarray:
    """Start offsets (no trailing entry) -> per-bag lengths.

    ``nn.EmbeddingBag(indices, offsets)`` as called at ``P/dlrm_s_pytorch.py:299-305``
    uses ``include_last_offset=False``: bag ``b`` covers ``[off[b], off[b+1])`` and the
    last bag runs to ``len(indices)``.  Same convention as the reference helper
    ``T/dlrm_s_pytorch.py:927-933`` (offsets -> lengths).
    """
    offsets = np.asarray(offsets, dtype=np.int64)
    ends = np.concatenate([offsets[1:], np.array([nnz], dtype=np.int64)])
    return ends - offsets


def offset2bag(offsets: np.ndarray, nnz: int) -> np.ndarray:
    """Bag id of every index position (ATen ``offset2bag`` of ``_embedding_bag``)."""
    return np.repeat(np.arange(len(offsets), dtype=np.int64), bag_lengths(offsets, nnz))


def linearize_indices_and_offsets(idx_list, off_list, rows_per_table):
    """Merged-module input layout used at ``P/dlrm_s_pytorch.py:908-915,1008-1014``.

    IPEX's source is not in the reference tree; SURVEY.md section 8b fixes the
    semantics from the call sites: concatenated indices, one CSR row-pointer array
    over all ``T*B`` bags (with the trailing total), and indices shifted by the
    cumulative row count of the preceding tables (global row ids).
    """
    idx_list = [np.asarray(i, dtype=np.int64) for i in idx_list]
    off_list = [np.asarray(o, dtype=np.int64) for o in off_list]
    nnz_base = np.concatenate([[0], np.cumsum([len(i) for i in idx_list])]).astype(np.int64)
    row_base = np.concatenate([[0], np.cumsum(rows_per_table)]).astype(np.int64)
    indices = np.concatenate(idx_list) if idx_list else np.zeros(0, np.int64)
    offsets = np.concatenate([o + nnz_base[t] for t, o in enumerate(off_list)] + [nnz_base[-1:]])
    iwro = np.concatenate([i + row_base[t] for t, i in enumerate(idx_list)]) if idx_list else indices
    return indices, offsets.astype(np.int64), iwro.astype(np.int64)


def sort_dedup(keys: np.ndarray):
    """Stable sort + run-length dedup of row ids.

    What the sparse gradient path needs before ``param.add_(sparse_grad)``
    (``torch.optim.SGD.step`` at ``P/dlrm_s_pytorch.py:1059``): equal to
    ``torch.sort(stable=True)`` / ``torch.unique(return_counts=True)`` /
    ``coalesce().indices()`` (KAT in SURVEY.md section 8c).
    Returns ``(sorted_keys, perm, unique_keys, counts, seg_offsets)``.
    """
    keys = np.asarray(keys)
    perm = np.argsort(keys, kind="stable").astype(np.int64)
    skeys = keys[perm]
    if len(keys) == 0:
        z = np.zeros(0, np.int64)
        return skeys, perm, z, z, np.zeros(1, np.int64)
    head = np.ones(len(keys), dtype=bool)
    head[1:] = skeys[1:] != skeys[:-1]
    starts = np.flatnonzero(head).astype(np.int64)
    seg = np.concatenate([starts, [len(keys)]]).astype(np.int64)
    return skeys, perm, skeys[starts], np.diff(seg), seg


def get_my_slice(n: int, world: int, rank: int) -> slice:
    """Contiguous table partition, ``P/extend_distributed.py:33-37``."""
    k, m = divmod(n, world)
    return slice(rank * k + min(rank, m), (rank + 1) * k + min(rank + 1, m), 1)


def get_split_lengths(n: int, world: int, rank: int):
    """``P/extend_distributed.py:39-49``: (my_len, splits or None when even)."""
    k, m = divmod(n, world)
    if m == 0:
        return k, None
    splits = [(k + 1) if i < m else k for i in range(world)]
    return splits[rank], splits


def lr_policy(step_count: int, base_lr: float, warmup: int, decay_start: int, decay_steps: int,
              last_lr: float | None = None) -> float:
    """``LRPolicyScheduler.get_lr`` ``P/dlrm_s_pytorch.py:146-167`` for one param group."""
    if step_count < warmup:
        return base_lr * (1.0 - (warmup - step_count) / warmup)
    if decay_start <= step_count < decay_start + decay_steps:
        scale = ((decay_steps - (step_count - decay_start)) / decay_steps) ** 2
        return max(0.0000001, base_lr * scale)
    if decay_steps > 0:
        return base_lr if last_lr is None else last_lr
    return base_lr


# --------------------------------------------------------------------------
# embedding bag forward / sparse backward + optimiser (tolerance stages)
# --------------------------------------------------------------------------


def embedding_bag_sum(W: np.ndarray, indices: np.ndarray, offsets: np.ndarray) -> np.ndarray:
    """``nn.EmbeddingBag(mode="sum")`` forward, ``P/dlrm_s_pytorch.py:299-305``.

    ``out[b] = sum_{j in bag b} W[indices[j]]``; an empty bag gives zeros.  Rows are
    added in index order with the accumulator in W's dtype (fp32), as ATen's CPU
    kernel does.
    """
    indices = np.asarray(indices, dtype=np.int64)
    nnz = len(indices)
    if nnz and (indices.min() < 0 or indices.max() >= W.shape[0]):
        raise IndexError("embedding index out of range")
    if len(offsets) and offsets[-1] > nnz:
        raise IndexError("offsets[-1] can not be greater than input's length")
    out = np.zeros((len(offsets), W.shape[1]), dtype=W.dtype)
    np.add.at(out, offset2bag(offsets, nnz), W[indices])
    return out


def embedding_bag_sparse_grad(indices, offsets, grad_out):
    """Autograd of the sum bag with ``sparse=True``: uncoalesced COO
    ``(indices, grad_out[offset2bag])`` (ATen ``_embedding_bag_sparse_backward``;
    triggered by ``E.backward()`` at ``P/dlrm_s_pytorch.py:1055``)."""
    indices = np.asarray(indices, dtype=np.int64)
    return indices, grad_out[offset2bag(offsets, len(indices))]


def sparse_sgd_update(W, indices, offsets, grad_out, lr):
    """``torch.optim.SGD.step`` on a sparse grad: ``W.add_(grad, alpha=-lr)``
    (``P/dlrm_s_pytorch.py:805,1059``).  Duplicated rows accumulate.  In place."""
    idx, vals = embedding_bag_sparse_grad(indices, offsets, grad_out)
    np.add.at(W, idx, (-np.float32(lr) * vals.astype(np.float32)).astype(W.dtype))
    return W


def sparse_adagrad_update(W, state_sum, indices, offsets, grad_out, lr, eps=1e-10):
    """``torch.optim.Adagrad`` sparse branch (no call site in the reference; BASELINE
    config 5): coalesce, ``state_sum += g^2``, ``W -= lr * g / (sqrt(state_sum)+eps)``."""
    idx, vals = embedding_bag_sparse_grad(indices, offsets, grad_out)
    _, _, uniq, _, _ = sort_dedup(idx)
    inv = np.searchsorted(uniq, idx)
    g = np.zeros((len(uniq), W.shape[1]), dtype=np.float32)
    np.add.at(g, inv, vals.astype(np.float32))
    state_sum[uniq] += g * g
    W[uniq] -= (np.float32(lr) * g / (np.sqrt(state_sum[uniq]) + np.float32(eps))).astype(W.dtype)
    return W, state_sum


# --------------------------------------------------------------------------
# pairwise dot-product interaction
# --------------------------------------------------------------------------


def tril_pairs(F: int, itself: bool = False):
    """Index lists of ``P/dlrm_s_pytorch.py:328-330`` (offset=0): strict lower
    triangle in row-major order (1,0),(2,0),(2,1),...; ``itself`` (``T/dlrm_s_pytorch.py:428-430``,
    offset=1) includes the diagonal: (0,0),(1,0),(1,1),(2,0),..."""
    o = 1 if itself else 0
    li = np.array([i for i in range(F) for j in range(i + o)], dtype=np.int64)
    lj = np.array([j for i in range(F) for j in range(i + o)], dtype=np.int64)
    return li, lj


def interact_fwd(x: np.ndarray, ly, itself: bool = False) -> np.ndarray:
    """``DLRM_Net.interact_features`` ``P/dlrm_s_pytorch.py:309-334`` (dot; self interaction only with
    ``itself``, ``T/dlrm_s_pytorch.py:407-433``): ``R = [x, (T T^t)[li, lj]]`` with ``T = stack([x] + ly)``."""
    T = np.stack([x] + list(ly), axis=1)  # [B, F, D]
    Z = np.matmul(T, T.transpose(0, 2, 1))
    li, lj = tril_pairs(T.shape[1], itself)
    return np.concatenate([x, Z[:, li, lj]], axis=1)


def interact_bwd(x: np.ndarray, ly, dR: np.ndarray, itself: bool = False):
    """Autograd of ``interact_fwd``: ``dT = (dZ + dZ^t) T`` with dZ scattered from
    ``dR[:, D:]``, plus the pass-through ``dR[:, :D]`` into x.  Returns (dx, [dly])."""
    T = np.stack([x] + list(ly), axis=1)
    B, F, D = T.shape
    li, lj = tril_pairs(F, itself)
    dZ = np.zeros((B, F, F), dtype=dR.dtype)
    dZ[:, li, lj] = dR[:, D:]
    dT = np.matmul(dZ + dZ.transpose(0, 2, 1), T)
    dx = dR[:, :D] + dT[:, 0]
    return dx, [dT[:, 1 + k] for k in range(F - 1)]


# ---------------------------------------------------------------- Criteo binary records
def criteo_transform(records: np.ndarray, max_ind_range: int, n_dense: int = 13):
    """``CriteoBinDataset.__getitem__`` + ``_transform_features``
    (``T/data_loader_terabyte.py:259-271`` and ``:83-102``, ``flag_input_torch_tensor=True``).

    ``records``: int32 ``[N, 1 + n_dense + T]`` = label, dense counts, categorical ids.
    Returns ``(X fp32 [N, n_dense], lS_o int64 [T, N], lS_i int64 [T, N], Y fp32 [N, 1])``:
    ``X = log(float32(x) + 1)``; ids ``% max_ind_range`` when it is > 0 with the sign of the
    divisor (torch ``%`` on an int tensor = floor-mod: never negative here); ``lS_o[t] = arange(N)``."""
    records = np.asarray(records, dtype=np.int32)
    y = records[:, 0].astype(np.float32).reshape(-1, 1)
    x_int = records[:, 1:1 + n_dense]
    x_cat = records[:, 1 + n_dense:]
    if max_ind_range > 0:
        x_cat = np.mod(x_cat, np.int32(max_ind_range))  # np.mod is floor-mod as well
    with np.errstate(divide="ignore", invalid="ignore"):
        X = np.log(x_int.astype(np.float32) + np.float32(1.0)).astype(np.float32)
    N, T = x_cat.shape
    lS_o = np.tile(np.arange(N, dtype=np.int64).reshape(1, -1), (T, 1))
    lS_i = np.ascontiguousarray(x_cat.astype(np.int64).T)
    return X, lS_o, lS_i, y


# ---------------------------------------------------------------- int8 tables for inference
def quantize_rowwise_int8(W: np.ndarray):
    """Symmetric int8 quantisation with one scale per ROW, zero point 0 -- the arithmetic of
    ``torch.ao.quantization.PerChannelMinMaxObserver(ch_axis=0, dtype=qint8, qscheme=per_channel_symmetric)``
    followed by ``torch.quantize_per_channel`` (the weight half of the reference's int8 QConfig,
    ``P/dlrm_s_pytorch.py:431-432,870-871``):
    ``scale = max(max(-min, max) / 127.5, eps_fp32)``, ``q = clamp(rint(w / scale), -128, 127)`` with fp32
    arithmetic and round-half-to-even.  Returns (q int8 [n, D], scale fp32 [n])."""
    W = np.asarray(W, dtype=np.float32)
    mn = np.minimum(W.min(axis=1), 0).astype(np.float32)
    mx = np.maximum(W.max(axis=1), 0).astype(np.float32)
    amax = np.maximum(-mn, mx).astype(np.float32)
    scale = np.maximum(amax / np.float32(127.5), np.float32(np.finfo(np.float32).eps)).astype(np.float32)
    inv = (np.float32(1.0) / scale).astype(np.float32)
    q = np.clip(np.rint((W * inv[:, None]).astype(np.float32)), -128, 127).astype(np.int8)
    return q, scale


def embedding_bag_sum_int8(q: np.ndarray, scale: np.ndarray, indices: np.ndarray, offsets: np.ndarray) -> np.ndarray:
    """Sum pooling over dequantised rows ``scale[r] * q[r]`` (fp32, index order); empty bag -> zeros."""
    deq = (q.astype(np.float32) * scale.astype(np.float32)[:, None]).astype(np.float32)
    return embedding_bag_sum(deq, indices, offsets)
