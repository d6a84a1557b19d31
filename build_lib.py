#!/usr/bin/env python
"""Build cloudtik_b200/lib/libb200dlrm.so (nvcc, sm_100a).  Stand-alone on purpose: the package itself refuses to
import while the library is missing, so the builder is loaded by path.

    python build_lib.py [--force] [--verbose]
"""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))


def load_builder():
    spec = importlib.util.spec_from_file_location("_b200_build", os.path.join(ROOT, "cloudtik_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def build(force=False, verbose=False):
    return load_builder().build(force=force, verbose=verbose)


if __name__ == "__main__":
    try:
        print("BUILD OK:", build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
    except Exception as e:  # keep the failure impossible to miss behind a pipe
        print(str(e)[-6000:])
        print("BUILD FAILED")
        sys.exit(1)
