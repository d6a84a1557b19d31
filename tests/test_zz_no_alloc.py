"""b200_sort_dedup_noalloc: the sort / dedup entry point for callers that cannot allocate.

Some users of this library are handed device pointers but cannot get device memory themselves:
their framework's allocator already owns all of HBM, or a cudaMalloc in their loop would stall the
whole device, or they are a thin C / cgo / JNI binding with no way to own a device buffer.  For
them the library carries its own fixed scratch (limit: B200_SORT_NOALLOC_MAX_KEYS keys) instead
of asking for a workspace.

This file is named so that pytest collects it last: the calls here share one scratch array inside
the library, and a failure here must not hide the other GPU tests from a `-x` run.  The test talks
to the C-ABI directly (ctypes) and compares with the existing workspace-taking entry point, bit for
bit (sorting and de-duplication are integer work), on two calls in a row -- the second call finds
the scratch dirty from the first, which must not matter.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_KEYS = 262144  # B200_SORT_NOALLOC_MAX_KEYS in include/b200dlrm.h


def _run(lib, entry, keys, vals, bound, ws=None):
    n = keys.numel()
    dev = keys.device
    out = {k: torch.full((n + (1 if k == "seg" else 0),), -1, dtype=torch.int32, device=dev)
           for k in ("skeys", "svals", "uniq", "seg")}
    nu = torch.full((1,), -1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    args = [keys.data_ptr(), vals.data_ptr(), n, bound, out["skeys"].data_ptr(), out["svals"].data_ptr(),
            out["uniq"].data_ptr(), out["seg"].data_ptr(), nu.data_ptr()]
    if ws is not None:
        args += [ws.data_ptr(), ws.numel()]
    rc = entry(*args, stream)
    assert rc == 0, lib.b200_last_error().decode()
    torch.cuda.synchronize()
    U = int(nu.item())
    return out["skeys"].cpu(), out["svals"].cpu(), out["uniq"][:U].cpu(), out["seg"][:U + 1].cpu(), U


def test_noalloc_matches_workspace_entry_point_twice():
    from cloudtik_b200 import _lib
    lib = _lib.lib
    rs = np.random.RandomState(2026)
    # first call: the largest size allowed, 4 radix passes; second: smaller, 2 passes, many duplicates
    for n, bound in ((MAX_KEYS, 187_767_399), (70_001, 5000)):
        keys = torch.tensor(rs.randint(0, bound, size=n).astype(np.uint32).view(np.int32)).cuda()
        vals = torch.arange(n, dtype=torch.int32, device="cuda")
        ws = torch.empty(int(lib.b200_sort_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
        want = _run(lib, lib.b200_sort_dedup, keys, vals, bound, ws)
        got = _run(lib, lib.b200_sort_dedup_noalloc, keys, vals, bound)
        assert got[4] == want[4]
        for g, w in zip(got[:4], want[:4]):
            assert torch.equal(g, w)
        # and both agree with the host answer
        k64 = keys.cpu().numpy().view(np.uint32).astype(np.int64)
        assert np.array_equal(got[0].numpy().view(np.uint32), np.sort(k64, kind="stable"))
        assert np.array_equal(got[1].numpy(), np.argsort(k64, kind="stable"))


def test_noalloc_rejects_oversize_without_launching():
    from cloudtik_b200 import _lib
    lib = _lib.lib
    one = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.b200_sort_dedup_noalloc(one.data_ptr(), one.data_ptr(), MAX_KEYS + 1, 10, one.data_ptr(), one.data_ptr(),
                                     one.data_ptr(), one.data_ptr(), one.data_ptr(),
                                     torch.cuda.current_stream().cuda_stream)
    assert rc == 2  # B200_EUNSUPPORTED
    msg = lib.b200_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(MAX_KEYS) in msg
