"""Build libb200dlrm.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m cloudtik_b200.build [--force] [--verbose]

The library has no torch dependency; it links the CUDA runtime statically so it loads
next to any torch build.  Objects are cached under cloudtik_b200/build/ by source mtime.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libb200dlrm.so")
SOURCES = ["api.cu", "emb_fwd.cu", "sort_dedup.cu", "emb_bwd.cu", "interact_simt.cu", "interact_tc.cu", "interact_tf32.cu", "interact_f32.cu", "criteo_input.cu", "mlp_eltwise.cu", "rowwise.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def _newest_header():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(os.path.dirname(HERE), "include", "b200dlrm.h"))
    return max(os.path.getmtime(h) for h in hs)


def _compile(src, force, verbose):
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    sp = os.path.join(CSRC, src)
    if (not force and os.path.exists(obj)
            and os.path.getmtime(obj) > max(os.path.getmtime(sp), _newest_header())):
        return obj, False, ""
    r = subprocess.run([NVCC] + FLAGS + ["-c", sp, "-o", obj], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    with open(obj + ".ptxas.log", "w") as f:
        f.write(r.stderr)
    return obj, True, r.stderr if verbose else ""


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIB_DIR, exist_ok=True)
    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        res = list(ex.map(lambda s: _compile(s, force, verbose), SOURCES))
    objs = [r[0] for r in res]
    for r in res:
        if r[2]:
            print(r[2])
    if any(r[1] for r in res) or not os.path.exists(LIB):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                   "-cudart", "static"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    try:
        print("BUILD OK:", build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
    except Exception as e:  # keep the failure impossible to miss behind a pipe
        print(str(e)[-6000:])
        print("BUILD FAILED")
        sys.exit(1)
