"""ctypes binding of libb200dlrm.so (the C-ABI declared in include/b200dlrm.h).

There is NO fallback: if the CUDA library is missing or a symbol cannot be bound, importing
fails loudly.  Build it with ``python build_lib.py`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200dlrm.so")

F32, BF16 = 0, 1
OPT_SGD, OPT_ADAGRAD = 0, 1

_c = ctypes
_p = _c.c_void_p
_i64 = _c.c_int64
_int = _c.c_int

# name -> (restype, argtypes); must list every symbol include/b200dlrm.h declares
SIGNATURES = {
    "b200_last_error": (_c.c_char_p, []),
    "b200_version": (_int, []),
    "b200_emb_fwd": (_int, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i64, _int, _int, _p, _p]),
    "b200_sort_workspace_bytes": (_c.c_size_t, [_i64]),
    "b200_sort_dedup": (_int, [_p, _p, _i64, _c.c_uint64, _p, _p, _p, _p, _p, _p, _c.c_size_t, _p]),
    "b200_sort_dedup_noalloc": (_int, [_p, _p, _i64, _c.c_uint64, _p, _p, _p, _p, _p, _p]),
    "b200_emb_bwd_workspace_bytes": (_c.c_size_t, [_i64, _i64]),
    "b200_emb_bwd_update": (_int, [_p, _p, _p, _p, _p, _p, _p, _i64, _p, _i64, _i64, _i64, _i64, _i64,
                                   _int, _int, _int, _c.c_float, _p, _c.c_float, _i64, _p,
                                   _c.c_size_t, _p, _p]),
    "b200_emb_bwd_prepare": (_int, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _p, _c.c_size_t, _p, _p]),
    "b200_emb_bwd_apply": (_int, [_p, _p, _p, _p, _i64, _p, _i64, _i64, _i64, _i64, _i64, _int, _int, _int,
                                  _c.c_float, _p, _c.c_float, _p, _c.c_size_t, _p]),
    "b200_emb_fwd_sharded": (_int, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _int, _int, _p, _p]),
    "b200_emb_bwd_apply_sharded": (_int, [_p, _p, _p, _p, _i64, _p, _i64, _i64, _i64, _i64, _i64, _i64, _int, _int, _int,
                                          _c.c_float, _p, _c.c_float, _p, _c.c_size_t, _p]),
    "b200_rowwise_filter_workspace_bytes": (_c.c_size_t, [_i64]),
    "b200_rowwise_filter": (_int, [_p, _p, _i64, _i64, _i64, _p, _i64, _p, _p, _p, _c.c_size_t, _p]),
    "b200_emb_fwd_sharded_from": (_int, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _int, _int, _int,
                                         _i64, _p, _p]),
    "b200_emb_fwd_sharded_add": (_int, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _int, _int, _p, _p]),
    "b200_emb_bwd_prepare_partial": (_int, [_p, _p, _p, _p, _i64, _p, _i64, _i64, _i64, _i64, _p, _c.c_size_t, _p, _p]),
    "b200_emb_bwd_coalesce": (_int, [_p, _p, _p, _p, _i64, _p, _i64, _i64, _i64, _i64, _i64, _int, _i64, _p, _p, _p,
                                     _p, _c.c_size_t, _p, _p]),
    "b200_emb_bwd_dense_grad": (_int, [_p, _i64, _p, _i64, _i64, _i64, _i64, _i64, _int, _p, _p, _c.c_size_t, _p]),
    "b200_interact_fwd_tf32": (_int, [_p, _p, _i64, _i64, _i64, _i64, _p]),
    "b200_interact_fwd_ex": (_int, [_p, _p, _i64, _i64, _i64, _i64, _int, _int, _p]),
    "b200_interact_bwd_ex": (_int, [_p, _p, _p, _i64, _i64, _i64, _i64, _int, _int, _p]),
    "b200_workspace_bytes": (_c.c_size_t, [_i64, _i64]),
    "b200_emb_bwd_sgd": (_int, [_p, _p, _p, _p, _p, _p, _i64, _p, _i64, _i64, _i64, _i64, _i64, _int, _int,
                                _c.c_float, _p, _i64, _p, _c.c_size_t, _p, _p]),
    "b200_emb_bwd_adagrad": (_int, [_p, _p, _p, _p, _p, _p, _p, _i64, _p, _i64, _i64, _i64, _i64, _i64, _int, _int,
                                    _c.c_float, _p, _c.c_float, _i64, _p, _c.c_size_t, _p, _p]),
    "b200_linearize": (_int, [_p, _p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _p]),
    "b200_interact_fwd": (_int, [_p, _p, _i64, _i64, _i64, _i64, _int, _p]),
    "b200_interact_bwd": (_int, [_p, _p, _p, _i64, _i64, _i64, _i64, _int, _p]),
    "b200_emb_fwd_q8": (_int, [_p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i64, _int, _p, _p]),
    "b200_relu_bwd_bias_workspace_bytes": (_c.c_size_t, [_i64, _i64]),
    "b200_relu_bwd_bias": (_int, [_p, _p, _p, _p, _i64, _i64, _int, _p, _c.c_size_t, _p]),
    "b200_criteo_transform": (_int, [_p, _i64, _int, _int, _i64, _int, _int, _i64, _i64, _p, _i64, _p, _p, _p, _p]),
}


class B200Error(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA library is not built. Run `python build_lib.py` at the repository root "
            "(needs nvcc); cloudtik_b200 has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc: int, what: str):
    if rc != 0:
        raise B200Error(f"{what} failed (status {rc}): {lib.b200_last_error().decode()}")
