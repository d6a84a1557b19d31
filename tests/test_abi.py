"""CPU-only: the C-ABI library loads and exports every symbol include/b200dlrm.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200dlrm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_expected_entry_points():
    names = _declared()
    for n in ("b200_emb_fwd", "b200_emb_bwd_update", "b200_sort_dedup", "b200_interact_fwd",
              "b200_interact_bwd", "b200_emb_bwd_workspace_bytes", "b200_sort_workspace_bytes"):
        assert n in names


def test_library_exports_every_declared_symbol():
    from cloudtik_b200 import _lib
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for n in _declared():
        assert hasattr(raw, n), f"{n} declared in include/b200dlrm.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in cloudtik_b200/_lib.py"
    assert _lib.lib.b200_version() >= 100


def test_host_side_argument_checks_need_no_gpu():
    from cloudtik_b200 import _lib
    # negative size -> EINVAL with a message, before any CUDA call
    rc = _lib.lib.b200_emb_fwd(None, None, None, None, None, 0, -1, 4, 8, 0, 0, 0, 0, None, None)
    assert rc == 1 and b"negative" in _lib.lib.b200_last_error()
    rc = _lib.lib.b200_sort_dedup(None, None, 5, 0, None, None, None, None, None, None, 0, None)
    assert rc == 2
    assert _lib.lib.b200_emb_bwd_workspace_bytes(1000, 128) > 6 * 4000
    # record wider than the staging buffer / table range outside the record
    rc = _lib.lib.b200_criteo_transform(None, 8, 100, 13, 0, 0, 26, 0, 8, None, 13, None, None, None, None)
    assert rc == 1 and b"record shape" in _lib.lib.b200_last_error()
    rc = _lib.lib.b200_criteo_transform(None, 8, 40, 13, 0, 20, 10, 0, 8, None, 13, None, None, None, None)
    assert rc == 1 and b"table range" in _lib.lib.b200_last_error()


def test_header_is_c99_and_library_links_from_plain_c(tmp_path):
    """include/b200dlrm.h compiles as C99 (-pedantic, warnings are errors) and tests/abi_client.c, linked against the
    library with no torch / Python in sight, gets the documented answers from the host-side checks."""
    import shutil
    import subprocess
    from cloudtik_b200 import _lib
    if shutil.which("gcc") is None:
        import pytest
        pytest.skip("no gcc")
    exe = tmp_path / "abi_client"
    libdir = os.path.dirname(_lib.LIB_PATH)
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "tests", "abi_client.c"), "-o", str(exe), "-L", libdir, "-lb200dlrm",
                        "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0 and "0 failure(s)" in r.stdout, r.stdout + r.stderr
