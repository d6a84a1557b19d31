"""B200-native DLRM training hot path behind the seams of CloudTik's AI runtime.

Everything here runs in ``lib/libb200dlrm.so`` (hand-written sm_100a kernels behind the C-ABI of
``include/b200dlrm.h``); there is no CPU path.  Importing the package without the built library fails right here,
with the command that builds it.
"""
import os as _os

_LIB = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "lib", "libb200dlrm.so")
if not _os.path.exists(_LIB) and _os.environ.get("CLOUDTIK_B200_BUILDING") != "1":
    raise ImportError(
        f"{_LIB} not found: cloudtik_b200 has no CPU fallback.  Build the CUDA library first: `python build_lib.py` "
        "at the repository root (needs nvcc; cross-compiles for sm_100a without a GPU).")
if _os.path.exists(_LIB):
    from . import _lib  # noqa: F401  (binds every symbol include/b200dlrm.h declares; fails loudly otherwise)
