"""Kernel micro-benchmarks (CUDA events, rotating index sets so rows are L2-cold).

    python tools/microbench.py [--rows 1000000] [--tables 26] [--dim 128] [--dtype fp32]
                               [--cases B:P,B:P,...] [--iters 20]

Prints one JSON line per (kernel, case): algorithmic bytes, time, GB/s,
fraction of the measured HBM peak (MEASURED_PEAKS.json) and of the nominal 8 TB/s.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cloudtik_b200 import ops  # noqa: E402


def peak_hbm():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


def time_ms(fn, iters, warmup=3):
    for _ in range(warmup):
        fn(0)
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for i, (a, b) in enumerate(evs):
        a.record()
        fn(i)
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2], ts[0]


C3 = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10, 2208, 11938, 155, 4,
      976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108, 36]


def run_cases(rows, D, dtype="fp32", act="", cases=((262144, 1), (8192, 32), (2048, 128), (32768, 1)), iters=20,
              n_sets=4, interact=True, dev="cuda"):
    """Generator of result dicts, one per (kernel, case).  ``rows``: rows per table; ``dtype``: table dtype;
    ``act``: activation / gradient dtype ("" = same as the tables)."""
    dt = torch.float32 if dtype == "fp32" else torch.bfloat16
    es = 4 if dtype == "fp32" else 2
    T = len(rows)
    F = T + 1
    peak, pk = peak_hbm()
    adt = dt if not act else (torch.float32 if act == "fp32" else torch.bfloat16)
    ea = 4 if adt == torch.float32 else 2
    tw = []
    for r in rows:
        w = torch.empty(r, D, dtype=dt, device=dev)
        w.uniform_(-0.5, 0.5)
        tw.append(w)
    ts = ops.TableSet(tw)
    rows_t = torch.tensor(rows, device=dev, dtype=torch.float64)
    g = torch.Generator(device=dev)
    g.manual_seed(1234)
    for B, P in cases:
        nnz_t = B * P
        sets = []
        for _ in range(n_sets):
            u01 = torch.rand(T, nnz_t, device=dev, generator=g, dtype=torch.float64)
            idx = torch.floor(u01 * rows_t[:, None]).to(torch.int64).reshape(-1)
            U = sum(int(torch.unique(idx[t * nnz_t:(t + 1) * nnz_t]).numel()) for t in range(T))
            sets.append((idx, U))
        off = torch.arange(0, T * B + 1, device=dev, dtype=torch.int64) * P
        buf = torch.empty(B, F, D, dtype=adt, device=dev)
        grad = torch.randn(B, F, D, device=dev).to(adt)
        ws = torch.empty(ops.emb_bwd_workspace_bytes(T * nnz_t, D), dtype=torch.uint8, device=dev)
        Uavg = sum(s[1] for s in sets) / len(sets)
        common = T * nnz_t * 8 + (T * B + 1) * 8
        fwd_bytes = Uavg * D * es + common + T * B * D * ea
        bwd_bytes = T * B * D * ea + common + 2 * Uavg * D * es

        def f_fwd(i):
            ops.emb_fwd(ts, sets[i % len(sets)][0], off, B, buf[:, 1:], D, F * D)

        def f_bwd(i):
            ops.emb_bwd_update(ts, sets[i % len(sets)][0], off, B, grad[:, 1:], D, F * D, lr=1e-6, ws=ws)

        def f_prep(i):
            ops.emb_bwd_prepare(ts, sets[i % len(sets)][0], off, B, ws)

        def f_apply(i):  # re-applies the last prepared index set (tiny lr keeps the weights sane)
            ops.emb_bwd_apply(ts, T * nnz_t, B, grad[:, 1:], D, F * D, 1e-6, ws)

        for name, fn, nbytes in (("emb_fwd", f_fwd, fwd_bytes), ("emb_bwd_sgd", f_bwd, bwd_bytes),
                                 ("emb_bwd_prepare", f_prep, common), ("emb_bwd_apply", f_apply, bwd_bytes)):
            med, best = time_ms(fn, iters)
            gbs = nbytes / med / 1e6
            yield {"kernel": name, "B": B, "P": P, "T": T, "D": D, "dtype": dtype, "ms": round(med, 4),
                   "ms_best": round(best, 4), "alg_GB": round(nbytes / 1e9, 4), "GBps": round(gbs, 1),
                   "frac_" + pk: round(gbs / peak, 3), "frac_8TBps": round(gbs / 8000, 3), "unique_rows": int(Uavg)}
        del buf, grad, ws, sets
    # interaction at the training shape
    for B in ((32768,) if interact else ()):
        Tt = torch.randn(B, F, D, device=dev).to(dt)
        R = torch.empty(B, D + F * (F - 1) // 2, dtype=dt, device=dev)
        dR = torch.randn_like(R)
        dT = torch.empty_like(Tt)
        fb = B * (F * D + D + F * (F - 1) // 2) * es
        bb = B * (2 * F * D + D + F * (F - 1) // 2) * es
        for name, fn, nbytes in (("interact_fwd", lambda i: ops.interact_fwd(Tt, R), fb),
                                 ("interact_bwd", lambda i: ops.interact_bwd(Tt, dR, dT), bb)):
            med, best = time_ms(fn, iters)
            gbs = nbytes / med / 1e6
            yield {"kernel": name, "B": B, "F": F, "D": D, "dtype": dtype, "ms": round(med, 4),
                   "alg_GB": round(nbytes / 1e9, 4), "GBps": round(gbs, 1), "frac_" + pk: round(gbs / peak, 3)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--tables", type=int, default=26)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--cases", default="262144:1,8192:32,2048:128,32768:1")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--sets", type=int, default=4)
    ap.add_argument("--c3", action="store_true", help="use the config-3 (Criteo-Terabyte) table sizes; fp32 tables")
    ap.add_argument("--c3-cap", type=int, default=0, help="cap config-3 table rows (0 = full, 96 GB)")
    ap.add_argument("--act", default="", help="activation / gradient dtype (default: same as --dtype)")
    ap.add_argument("--no-interact", action="store_true")
    a = ap.parse_args()
    rows = [min(r, a.c3_cap) if a.c3_cap else r for r in C3] if a.c3 else [a.rows] * a.tables
    cases = [tuple(int(v) for v in c.split(":")) for c in a.cases.split(",")]
    for r in run_cases(rows, a.dim, a.dtype, a.act, cases, a.iters, a.sets, not a.no_interact):
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
