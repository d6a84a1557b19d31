"""Forward-only (serving) throughput of the config-3 DLRM on one B200: fp32 tables vs int8 tables.

    python tools/infer_bench.py [--batch 32768] [--steps 50] [--rows-cap 0]

The counterpart of the reference's ``--inference-only`` / ``run_throughput_benchmark`` runs
(``QS/scripts/recommendation/pytorch/dlrm/inference/cpu/inference_performance.sh:76-85``): bf16 MLPs and
interaction, whole forward replayed as one CUDA graph, inputs resident in HBM, CUDA-event timing.  Prints one
JSON line per table format with samples/s, ms per batch and the pooled-embedding kernel's time and
algorithmic GB/s (CUDA events around single launches in an eager pass).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from bench import C3_ROWS, D, LN_BOT, LN_TOP  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32768)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--rows-cap", type=int, default=0)
    a = ap.parse_args()
    from cloudtik_b200 import ops
    from cloudtik_b200.dlrm import DLRM_Net
    from cloudtik_b200.inference import GraphedInference
    from cloudtik_b200.quantized import quantize_embeddings_int8

    dev = torch.device("cuda", 0)
    rows = [min(r, a.rows_cap) if a.rows_cap else r for r in C3_ROWS]
    np.random.seed(727)
    net = DLRM_Net(D, np.array(rows), np.array(LN_BOT), np.array(LN_TOP), sigmoid_top=len(LN_TOP) - 2,
                   np_init_emb_weight=False, device=dev, bf16=True)
    T, B = len(rows), a.batch
    rs = np.random.RandomState(1)
    sets = []
    for _ in range(4):
        X = torch.tensor(rs.rand(B, 13).astype(np.float32), device=dev)
        idx = np.stack([np.floor(rs.rand(B) * n).astype(np.int64) for n in rows])
        uniq = sum(int(np.unique(r).size) for r in idx)
        sets.append((X, torch.tensor(idx.reshape(-1), device=dev), torch.arange(T * B + 1, device=dev), uniq))
    U = statistics.mean(s[3] for s in sets)

    def run(tag, wbytes, fwd_call):
        gi = GraphedInference(net, sets[0][:3])
        for i in range(5):
            gi(*sets[i % 4][:3])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.steps):
            gi(*sets[i % 4][:3])
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        # the pooled-embedding kernel alone
        out = torch.empty(B, T + 1, D, dtype=torch.bfloat16, device=dev)
        ts = []
        for i in range(12):
            torch.cuda._sleep(300_000)
            x, y = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            x.record()
            fwd_call(sets[i % 4][1], sets[i % 4][2], out)
            y.record()
            torch.cuda.synchronize()
            ts.append(x.elapsed_time(y))
        k_ms = statistics.median(ts[2:])
        alg = U * D * wbytes + (U * 4 if wbytes == 1 else 0) + T * B * 8 + (T * B + 1) * 8 + T * B * D * 2
        print(json.dumps({"tables": tag, "batch": B, "ms_per_batch": round(ms, 4),
                          "samples_per_s": round(B / ms * 1e3, 1), "emb_fwd_ms": round(k_ms, 4),
                          "emb_fwd_alg_GB": round(alg / 1e9, 4), "emb_fwd_GBps": round(alg / k_ms / 1e6, 1),
                          "table_bytes_GB": round(sum(rows) * D * wbytes / 1e9, 1)}), flush=True)

    emb = net.emb_l
    run("fp32", 4, lambda i, o, out: ops.emb_fwd(emb.tables, i, o, B, out[:, 1:], D, (T + 1) * D))
    quantize_embeddings_int8(net)
    q = net.emb_l
    run("int8 + fp32 row scale", 1, lambda i, o, out: ops.emb_fwd_q8(q.tables, i, o, B, out[:, 1:], D, (T + 1) * D))


if __name__ == "__main__":
    main()
