#!/usr/bin/env python
"""DLRM training-step benchmark (BASELINE.json metric: DLRM train samples/sec; EmbeddingBag HBM GB/s).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--dtype bf16|fp32]

Workload at N=1: BASELINE config 3 -- DLRM Criteo-Terabyte shape: 26 tables (MLPerf cardinalities
under the 40M cap, 187.8M rows, 96 GB fp32), D=128, bottom 13-512-256-128, top 479-1024-1024-512-256-1,
batch 32768, one-hot, SGD, BCE; synthetic uniform indices, random-init weights.
At N>1 (config 4 family): tables sharded table-wise over ranks (ext_dist.get_my_slice), pooled
all-to-all + DDP allreduce over NCCL, 32768 samples PER GPU (weak scaling; N=2 is exactly the
64K global batch of config 4).

A "step" = forward + BCE loss + backward + update (fused sparse SGD on the tables inside backward,
torch SGD on the MLPs).  `value` = samples/s with inputs resident in HBM; `e2e` = the same through
the public module API with inputs in pinned HOST memory (H2D every step) and the loss read back
(D2H) every step.  Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# MLPerf Criteo-1TB cardinalities under --max-ind-range=40000000
C3_ROWS = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10, 2208,
           11938, 155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108, 36]
D = 128
LN_BOT = [13, 512, 256, D]
LN_TOP = [D + 27 * 26 // 2, 1024, 1024, 512, 256, 1]
LR = 1.0  # QS/scripts/recommendation/pytorch/dlrm/training/cpu/training.sh:83-93


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-fp32", action="store_true", help="skip the extra fp32 leg of the 1-GPU line")
    ap.add_argument("--batch", type=int, default=32768, help="samples per GPU (weak scaling: fixed as N grows)")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="strong scaling: fixed GLOBAL batch split over the GPUs (BASELINE configs[3]: 65536); "
                         "overrides --batch and reports \"scaling\": \"strong\"")
    ap.add_argument("--chunks", type=int, default=0,
                    help="multi-GPU peer-memory path: micro-batches per step (default 1); with several, one chunk's "
                         "NVLink exchange runs under another chunk's compute (measured slower at 2 GPUs: the step's "
                         "many small kernels do not shrink with the chunk)")
    ap.add_argument("--no-parity", action="store_true", help="multi-GPU: skip the pre-flight parity check")
    ap.add_argument("--nccl-tail", action="store_true",
                    help="multi-GPU: NCCL instead of the symmetric-memory two-shot all-reduce for the end-of-step bucket")
    ap.add_argument("--sparse-dense-boundary", type=int, default=2048,
                    help="multi-GPU: tables with fewer rows are replicated on every rank instead of sharded (the "
                         "reference training variant's flag and default, T/dlrm_s_pytorch.py:266-269); 0 = shard all")
    ap.add_argument("--rows-cap", type=int, default=0, help="debug: cap table rows (0 = full config 3)")
    ap.add_argument("--sets", type=int, default=4, help="distinct synthetic batches rotated")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="run the step eagerly instead of as one CUDA graph")
    ap.add_argument("--no-p2p", action="store_true", help="multi-GPU: NCCL all-to-all instead of the fused peer-memory exchange")
    ap.add_argument("--workload", default="c3", choices=["c3", "c5"],
                    help="c3 = Criteo-Terabyte one-hot step (the headline line); c5 = BASELINE configs[4]: multi-hot "
                         "bags, row-wise sharded hot tables, fused Adagrad, bf16 activations (eager, N >= 1)")
    ap.add_argument("--pooling", type=int, default=20, help="c5: indices per bag")
    ap.add_argument("--no-c5", action="store_true",
                    help="skip the BASELINE configs[4] (multi-hot, row-wise sharded, fused Adagrad) block appended to the line")
    ap.add_argument("--no-microbench", action="store_true",
                    help="skip the BASELINE configs[1] EmbeddingBag micro-benchmark appended to the N=1 line")
    ap.add_argument("--timeline", default="", help="after the timed legs: profile 3 more steps with torch.profiler and "
                    "write every rank's kernel list (name, stream, start, duration) to <path>.rank<r>.json")
    ap.add_argument("--cpu-rows-cap", type=int, default=200000)
    ap.add_argument("--cpu-steps", type=int, default=24, help="timed steps of the cpu_baseline leg (~10 s of CPU work)")
    return ap.parse_args()


def table_rows(cap):
    return [min(r, cap) if cap else r for r in C3_ROWS]


# ------------------------------------------------------------------ CPU reference arm / baseline
def cpu_reference_run(batch, rows_cap, steps, warmup, bf16, threads=None):
    """The reference's own CPU implementation of the step, as restated in oracle/dlrm_port.py
    (the reference tree itself is not on the GPU box).  All host threads torch can use."""
    from oracle.dlrm_port import DLRMPort, train_step, uniform_batch
    if threads:
        torch.set_num_threads(threads)
    rows = table_rows(rows_cap)
    np.random.seed(727)
    net = DLRMPort(D, rows, LN_BOT, LN_TOP)
    opt = torch.optim.SGD(net.parameters(), lr=LR)
    rs = np.random.RandomState(727)
    X, lS_o, lS_i, T = uniform_batch(rs, batch, rows, pooling=1)
    X, T = torch.tensor(X), torch.tensor(T)
    lS_o = [torch.tensor(o) for o in lS_o]
    lS_i = [torch.tensor(i) for i in lS_i]
    for _ in range(warmup):
        train_step(net, opt, X, lS_o, lS_i, T, bf16=bf16)
    t0 = time.time()
    for _ in range(steps):
        train_step(net, opt, X, lS_o, lS_i, T, bf16=bf16)
    dt = (time.time() - t0) / max(steps, 1)
    return {"value": batch / dt, "unit": "samples/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"config-3 architecture, batch {batch}, table rows capped at {rows_cap}, "
                      f"{steps} steps after {warmup} warm-up, torch {torch.__version__} CPU, "
                      f"{'bf16 autocast' if bf16 else 'fp32'}",
            "ms_per_step": dt * 1e3}


def host_threads():
    """Every host core this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers: the CPU
    arm must not inherit that (the reference's own launcher gives a rank all cores of its socket,
    QS/scripts/recommendation/pytorch/dlrm/training/cpu/distribute_training.sh:73-105)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def batch_sizes(a, world):
    """(samples per GPU, global batch, scaling) for this run."""
    if a.global_batch:
        if a.global_batch % world:
            sys.exit("--global-batch must be a multiple of the number of GPUs")
        return a.global_batch // world, a.global_batch, "strong"
    return a.batch, a.batch * world, "weak"


def run_reference_arm(a):
    """The reference's CPU implementation of the step (pinned port, oracle/dlrm_port.py) on ALL host threads, on
    the SAME global batch the B200 arm processes per step at this N (one process: the CPU arm has no ranks).  The
    96 GB of tables do not fit the host, so the rows are capped (same architecture, same batch, `same_config`
    false for that reason only; a smaller table is cache-friendlier, i.e. the cap favours the CPU)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    _, B_glob, scaling = batch_sizes(a, max(world, a.gpus, 1))
    # ~11 us per sample on 16 threads: bound the whole run to about a minute
    est = B_glob * 11e-6
    steps = max(1, min(a.steps, 30, int(60.0 / est)))
    warm = max(1, min(max(a.warmup, 3), 3, int(15.0 / est)))
    r = cpu_reference_run(B_glob, a.cpu_rows_cap, steps, warm, a.dtype == "bf16", threads=host_threads())
    line = {"impl": "reference", "metric": "dlrm_train_samples_per_sec", "value": r["value"], "unit": "samples/s",
            "n_gpus": a.gpus, "steps": steps, "warmup": warm, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": a.dtype, "data": "synthetic",
            "config": {"workload": "dlrm_criteo_terabyte_c3", "global_batch": B_glob, "dim": D,
                       "tables": len(C3_ROWS), "cpu_table_rows_cap": a.cpu_rows_cap, "same_config": False,
                       "same_config_note": f"same architecture and global batch; table rows capped at "
                                           f"{a.cpu_rows_cap} (96 GB of tables do not fit host memory)",
                       "host_threads": r["cores"], "processes": 1},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "20", "-i", str(index)], stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in out.strip().splitlines():
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------ B200 arm
def peak_hbm():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """dram bytes per launch from the committed ncu capture (profiles/roofline_traffic.json), or None."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json"))).get(kernel)
    except Exception:
        return None


def make_batches(a, rows_local, B_loc, B_glob, dev, rank, n_chunks=1, rows_repl=()):
    """Synthetic Criteo-shaped inputs: dense U[0,1), targets round(U), one index per
    (sample, table) uniform over the table's rows.  Generated on the host with numpy, kept both in
    pinned host memory (e2e leg) and on the device (HBM-resident leg).
    A batch is the flat tuple (X, Y, indices_0, offsets_0[, indices_1, offsets_1, ...]): one pair of linearised
    bags per micro-batch, micro-batch c = the c-th slice of every rank's samples (dlrm.train_step_chunked)."""
    rs = np.random.RandomState(727 + 1000 * rank)
    T = len(rows_local)
    sets = []
    Bc = B_glob // n_chunks
    for _ in range(a.sets):
        X = rs.rand(B_loc, 13).astype(np.float32)
        Y = np.round(rs.rand(B_loc, 1)).astype(np.float32)
        idx = np.empty((T, B_glob), dtype=np.int64)
        for t, n in enumerate(rows_local):
            idx[t] = np.floor(rs.rand(B_glob) * n).astype(np.int64)
        flat = [X, Y]
        for c in range(n_chunks):
            flat.append(np.ascontiguousarray(idx[:, c * Bc:(c + 1) * Bc]).reshape(-1))
            flat.append(np.arange(T * Bc + 1, dtype=np.int64))
        if len(rows_repl):  # replicated small tables: this rank's LOCAL bags, as the last pair of the tuple
            ridx = np.stack([np.floor(rs.rand(B_loc) * n).astype(np.int64) for n in rows_repl])
            flat.append(ridx.reshape(-1))
            flat.append(np.arange(len(rows_repl) * B_loc + 1, dtype=np.int64))
        host = [torch.from_numpy(v).pin_memory() for v in flat]
        # the same batch as RAW Criteo records (int32 label, 13 counts, T ids: the CriteoBinDataset format);
        # the counts are chosen so that log(count + 1) reproduces a dense feature of the same range
        host_rec = None
        if B_loc == B_glob:  # single GPU only (a rank of a sharded job would need every table's ids)
            rec = np.empty((B_glob, 14 + T), dtype=np.int32)
            rec[:, 0] = Y[:, 0]
            rec[:, 1:14] = np.floor(np.expm1(4.0 * X))
            rec[:, 14:] = idx.T
            host_rec = torch.from_numpy(rec).pin_memory()
        devt = [h.to(dev, non_blocking=True) for h in host]
        uniq = [int(np.unique(idx[t]).size) for t in range(T)]
        sets.append({"host": host, "dev": devt, "uniq": uniq, "host_rec": host_rec})
    torch.cuda.synchronize()
    return sets


def fp32_leg(a, rows, sets, dev):
    """The reference's DEFAULT precision (fp32 everywhere: P/dlrm_s_pytorch.py:707,1029 turn bf16 autocast on only with
    --bf16) on the same config-3 workload and the same device batches: fp32 MLPs (cuBLAS), fp32 pooled rows, fp32
    interaction (CUDA-core kernels; the 3xTF32 tcgen05 forward is opt-in), fused fp32 update; one CUDA graph per step.
    Runs last (the bf16 model is gone by then) and is contained: a failure becomes an "error" entry."""
    import gc

    from cloudtik_b200.dlrm import DLRM_Net
    from cloudtik_b200.graph import GraphedTrainStep
    try:
        gc.collect()
        torch.cuda.empty_cache()
        np.random.seed(727)
        torch.manual_seed(727)
        net = DLRM_Net(D, np.array(rows), np.array(LN_BOT), np.array(LN_TOP), sigmoid_top=len(LN_TOP) - 2,
                       np_init_emb_weight=False, device=dev, lr=LR, bf16=False)
        opt = torch.optim.SGD([p for p in net.parameters() if p.requires_grad], lr=LR)
        gs = GraphedTrainStep(net, opt, sets[0]["dev"])
        last = None
        for i in range(max(a.warmup, 3)):
            last = gs.step(*sets[i % len(sets)]["dev"])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.steps):
            last = gs.step(*sets[i % len(sets)]["dev"])
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        out = {"value": round(a.batch / (ms / 1e3), 1), "unit": "samples/s", "ms_per_step": round(ms, 4),
               "dtype": "fp32", "steps": a.steps, "final_loss": float(last.float().cpu()),
               "interaction": "fp32 CUDA-core kernels" + (" / 3xTF32 tcgen05 forward" if os.environ.get(
                   "B200_INTERACT_TF32") == "1" else ""),
               "mlp": "cuBLAS fp32 on CUDA cores (TF32 off: the reference's fp32 results are the 1e-5 parity target)",
               "note": "same workload, batches and graph replay as the headline leg, inputs resident in HBM"}
        del gs, net, opt
        return out
    except Exception as e:  # noqa: BLE001 -- reported in the line, the headline legs are already measured
        return {"error": f"{type(e).__name__}: {e}"[:300]}


def embeddingbag_microbench(net, gs):
    """BASELINE.json configs[1] (the configuration the metric's "EmbeddingBag HBM GB/s vs ~8 TB/s" half is
    quoted on): 26 tables x 1 M rows, dim 128, pooling 1 / 32 / 128, fp32 and bf16 tables, forward and the
    fused backward+update.  CUDA events around single launches, rotating index sets (rows L2-cold),
    algorithmic bytes of SURVEY.md section 8d.  Runs after the timed regions; the step's model is released first."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gc

    import microbench as mb
    del gs
    net.emb_l = None  # 96 GB of tables are not needed any more
    gc.collect()
    torch.cuda.empty_cache()
    out = []
    for dtype in ("fp32", "bf16"):
        for r in mb.run_cases([1_000_000] * 26, D, dtype, "", ((262144, 1), (8192, 32), (2048, 128)), iters=10,
                              n_sets=3, interact=False):
            if r["kernel"] in ("emb_fwd", "emb_bwd_apply", "emb_bwd_sgd"):
                out.append({k: r[k] for k in ("kernel", "dtype", "B", "P", "ms", "GBps", "frac_8TBps")}
                           | {"frac_measured": r.get("frac_measured")})
        gc.collect()
        torch.cuda.empty_cache()
    return {"config": "26 tables x 1M rows, dim 128, pooling 1/32/128 (BASELINE configs[1])",
            "kernels": "emb_fwd = pooled forward; emb_bwd_apply = segment-reduce + row update; emb_bwd_sgd = "
                       "index sort/dedup + emb_bwd_apply back to back on one stream (in the training step the index stage "
                       "runs on a side stream under the dense forward/backward)", "results": out}


def _dev(a_, b_, rtol, atol):
    """allclose-style deviation, scaled so that `<= rtol` means |a - b| <= atol + rtol * |b| everywhere."""
    return float(((a_ - b_).abs() / (atol + rtol * b_.abs())).max()) * rtol


def parity_preflight(world, rank, dev, use_p2p, n_chunks, use_graph, boundary=0, steps=3):
    """Multi-GPU parity, checked in the same process right before the timed run: a small-table model of the bench's
    architecture takes `steps` training steps through EXACTLY the sharded path the bench times (table-wise shards,
    peer-memory exchange or NCCL all-to-all, micro-batches, CUDA graph, flat MLP all-reduce, lr / world on the
    tables), and every rank also runs the single-process model (DLRM_Net(distributed=False), the path the one-GPU
    tests pin to the oracle) on the GLOBAL batch with the same initial weights.  The single-process global-batch run
    is the multi-GPU oracle of the reference (T/dlrm_s_pytorch.py:483-522, lr/world at
    :1020-1031).  Compared: the mean-of-ranks loss of every step and, after the last step, the rows of the tables
    this rank owns (sharded and replicated) and the MLP weights, with the tolerances of tests/dist_parity.py:
    |a - b| <= atol + tol * |b|, atol 1e-6 (fp32) / 2e-3 (bf16) for table rows, 10 x tol and 5e-6 / 2e-3 for the
    MLPs; the reported figures are scaled so that `<= tol` is a pass."""
    from cloudtik_b200 import ext_dist
    from cloudtik_b200.dlrm import DLRM_Net, enable_peer_exchange, train_step_chunked
    from cloudtik_b200.graph import GraphedTrainStep
    rows = [min(r, 3000) for r in C3_ROWS]
    T = len(rows)
    lB = 1024
    Bg = lB * world
    lr = 0.5
    out = {"steps": steps, "tables": T, "rows_cap": 3000, "global_batch": Bg, "micro_batches": n_chunks,
           "exchange": "peer-memory" if use_p2p else "nccl all-to-all", "graph": bool(use_graph),
           "sparse_dense_boundary": boundary,
           "oracle": "single-process cloudtik_b200 run on the global batch (every rank)"}
    ok_all = True
    for tag, bf16, tol in (("fp32", False, 1e-5), ("bf16", True, 1e-2)):
        g = torch.Generator(device="cpu").manual_seed(4242)
        W0 = [(torch.rand(n, D, generator=g) * 2 - 1) * (1.0 / n) ** 0.5 for n in rows]
        np.random.seed(99)
        ref = DLRM_Net(D, np.array(rows), np.array(LN_BOT), np.array(LN_TOP), sigmoid_top=len(LN_TOP) - 2,
                       device=dev, lr=lr, bf16=bf16, distributed=False, sparse_dense_boundary=boundary,
                       dense_tables_first=True)
        np.random.seed(99)
        net = DLRM_Net(D, np.array(rows), np.array(LN_BOT), np.array(LN_TOP), sigmoid_top=len(LN_TOP) - 2,
                       device=dev, lr=lr / world, bf16=bf16, sparse_dense_boundary=boundary)
        order = ref.table_order  # table behind each feature slot (small replicated tables first when boundary > 0)

        def reset_tables():  # same MLPs by the same numpy stream; tables from the common draw
            with torch.no_grad():
                for k, w in enumerate(ref.emb_l.weights):
                    w.copy_(W0[order[k]])
                for k, t in enumerate(net.local_ln_emb):
                    net.emb_l.weights[k].copy_(W0[t])
                if net.emb_dense is not None:
                    for p_, t in zip(net.emb_dense.params(), net.ln_emb_dense):
                        p_.copy_(W0[t])
        reset_tables()
        if use_p2p:
            enable_peer_exchange(net, lB, n_chunks)
        nc = n_chunks if use_p2p else 1
        dense_r = [p for p in ref.parameters() if p.requires_grad]
        dense_n = [p for p in net.parameters() if p.requires_grad]
        opt_r, opt_n = torch.optim.SGD(dense_r, lr=lr), torch.optim.SGD(dense_n, lr=lr)
        sync = ext_dist.DenseGradAllreduce(dense_n, early=list(net.top_l.parameters()), symmetric_tail=use_p2p)
        rs = np.random.RandomState(5)
        batches = []
        for _ in range(steps):
            X = rs.rand(Bg, 13).astype(np.float32)
            Y = np.round(rs.rand(Bg, 1)).astype(np.float32)
            idx = np.stack([np.floor(rs.rand(Bg) * n).astype(np.int64) for n in rows])  # [T, Bg]
            batches.append((X, Y, idx))

        def shard(X, Y, idx):
            """This rank's view: its samples, and its tables' ids for every rank's samples, micro-batch by micro-batch."""
            mine = [torch.from_numpy(X[rank * lB:(rank + 1) * lB]).to(dev), torch.from_numpy(Y[rank * lB:(rank + 1) * lB]).to(dev)]
            loc = idx[net.local_ln_emb]                       # [T_loc, Bg], sample b belongs to rank b // lB
            per = lB // nc
            v = loc.reshape(len(loc), world, nc, per)
            for c in range(nc):
                mine.append(torch.from_numpy(np.ascontiguousarray(v[:, :, c]).reshape(-1)).to(dev))
                mine.append(torch.arange(len(loc) * world * per + 1, dtype=torch.int64, device=dev))
            if net.emb_dense is not None:
                rep = idx[net.ln_emb_dense][:, rank * lB:(rank + 1) * lB]
                mine.append(torch.from_numpy(np.ascontiguousarray(rep).reshape(-1)).to(dev))
                mine.append(torch.arange(rep.size + 1, dtype=torch.int64, device=dev))
            return mine

        gs = GraphedTrainStep(net, opt_n, shard(*batches[0]), post_backward=sync) if use_graph else None
        mlp_n = [p for n_, p in net.named_parameters() if p.requires_grad and not n_.startswith("emb_dense")]
        if gs is not None:  # the capture's warm-up steps trained on batch 0: start again from the common weights
            reset_tables()
            with torch.no_grad():
                for pn, pr in zip(mlp_n, dense_r):
                    pn.copy_(pr)
        worst_loss = 0.0
        for X, Y, idx in batches:
            b = shard(X, Y, idx)
            E = gs.step(*b) if gs is not None else train_step_chunked(net, opt_n, b, sync)
            E = E.clone()
            torch.distributed.all_reduce(E)
            got = float(E.item()) / world
            gb = [torch.from_numpy(X).to(dev), torch.from_numpy(Y).to(dev),
                  torch.from_numpy(np.ascontiguousarray(idx[order]).reshape(-1)).to(dev),
                  torch.arange(T * Bg + 1, dtype=torch.int64, device=dev)]
            want = float(train_step_chunked(ref, opt_r, gb).item())
            worst_loss = max(worst_loss, abs(got - want) / abs(want))
        worst_rows = 0.0
        slot = {t: k for k, t in enumerate(order)}  # table id -> position in the single-process model
        atol_w, atol_m = (2e-3, 2e-3) if bf16 else (1e-6, 5e-6)
        mine = [(net.emb_l.weights[k], t) for k, t in enumerate(net.local_ln_emb)]
        if net.emb_dense is not None:
            mine += [(p_.data, t) for p_, t in zip(net.emb_dense.params(), net.ln_emb_dense)]
        for w_, t in mine:
            worst_rows = max(worst_rows, _dev(w_.float(), ref.emb_l.weights[slot[t]].float(), tol, atol_w))
        worst_mlp = max(_dev(pn.data, pr.data, 10 * tol, atol_m) / 10 for pn, pr in zip(mlp_n, dense_r))
        w = torch.tensor([worst_loss, worst_rows, worst_mlp], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(w, op=torch.distributed.ReduceOp.MAX)
        worst_loss, worst_rows, worst_mlp = (float(x) for x in w)
        ok = worst_loss <= tol and worst_rows <= tol and worst_mlp <= tol
        ok_all = ok_all and ok
        out[tag] = {"max_rel_loss": worst_loss, "max_rel_rows": worst_rows, "max_rel_mlp": worst_mlp, "tol": tol,
                    "ok": ok}
        del net, ref, gs
        torch.cuda.synchronize()
        torch.distributed.barrier()
    out["ok"] = ok_all
    return out


def write_timeline(path, rank, step_fn, barrier, steps=3):
    """Kernel timeline of `steps` more steps of the timed loop on THIS rank (torch.profiler / CUPTI sees the kernels
    of a replayed graph too).  One JSON per rank; tools/timeline_summary.py turns them into the per-rank table."""
    from torch.profiler import ProfilerActivity, profile
    barrier()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(steps):
            step_fn(i)
        torch.cuda.synchronize()
    ev = []
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA:
            ev.append({"name": e.name, "ts_us": e.time_range.start, "dur_us": e.time_range.elapsed_us(),
                       "stream": getattr(e, "device_resource_id", None)})
    ev.sort(key=lambda x: x["ts_us"])
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(f"{path}.rank{rank}.json", "w") as f:
        json.dump({"rank": rank, "steps": steps, "events": ev}, f)
    barrier()


def run_b200_arm(a):
    from cloudtik_b200 import ext_dist
    from cloudtik_b200.dlrm import DLRM_Net, enable_peer_exchange

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        ext_dist.init_distributed(rank=rank, size=world, backend="nccl")
    bf16 = a.dtype == "bf16"
    rows = table_rows(a.rows_cap)
    B_loc, B_glob, scaling = batch_sizes(a, world)
    use_p2p = world > 1 and not a.no_p2p
    n_chunks = a.chunks if (a.chunks > 0 and use_p2p) else 1
    boundary = a.sparse_dense_boundary if use_p2p and n_chunks == 1 else 0
    parity = None
    if world > 1 and not a.no_parity:
        parity = parity_preflight(world, rank, dev, use_p2p, n_chunks, not a.no_graph, boundary)

    np.random.seed(727)
    torch.manual_seed(727 + rank)
    # sharded tables see the sum of every rank's local-mean gradient: lr/world (T/dlrm_s_pytorch.py:1020-1031)
    net = DLRM_Net(D, np.array(rows), np.array(LN_BOT), np.array(LN_TOP), sigmoid_top=len(LN_TOP) - 2,
                   np_init_emb_weight=False, device=dev, lr=LR / world, bf16=bf16, sparse_dense_boundary=boundary)
    local_ids = list(range(len(rows))) if world == 1 else net.local_ln_emb
    rows_local = [rows[i] for i in local_ids]
    rows_repl = [rows[i] for i in net.ln_emb_dense] if world > 1 else []
    if use_p2p:
        enable_peer_exchange(net, B_loc, n_chunks)
    dense_params = [p for p in net.parameters() if p.requires_grad]
    opt = torch.optim.SGD(dense_params, lr=LR)
    if bf16:  # bf16 copies of the MLP parameters refreshed by one multi-tensor copy after opt.step()
        from cloudtik_b200.mlp import enable_bf16_shadows
        enable_bf16_shadows([net.bot_l, net.top_l], opt)
    # MLPs are data-parallel: one flat all-reduce of their 2.4 M gradients per step (the fused bf16 MLP
    # path bypasses DDP.forward, so the averaging the reference gets from DDP is done explicitly)
    # (the top MLP's share is reduced from inside backward, under the interaction backward and the table update)
    sync_dense = (ext_dist.DenseGradAllreduce(dense_params, early=list(net.top_l.parameters()),
                                              symmetric_tail=use_p2p and not a.nccl_tail)
                  if world > 1 else (lambda: None))
    sets = make_batches(a, rows_local, B_loc, B_glob, dev, rank, n_chunks, rows_repl)
    emb = net.emb_l
    from cloudtik_b200.dlrm import train_step_chunked
    post = sync_dense if world > 1 else None

    def step_dev(i):
        return train_step_chunked(net, opt, sets[i % len(sets)]["dev"], post)

    stage = [torch.empty_like(t) for t in sets[0]["dev"]]

    def step_e2e(i):
        host = sets[i % len(sets)]["host"]
        for d, h in zip(stage, host):
            d.copy_(h, non_blocking=True)
        E = train_step_chunked(net, opt, stage, post)
        return float(E.cpu())  # D2H read of the step's loss (reference: P/dlrm_s_pytorch.py:1048)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        for i in range(steps):
            last = fn(i)
        if hasattr(fn, "flush"):
            fn.flush()  # losses still in flight are read on the host before the clock stops
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        return float(ms.item()), last

    for i in range(max(a.warmup, 3)):
        step_dev(i)
    # per-kernel CUDA-event timing of the hand-written embedding kernels (roofline): an eager pass of
    # the same K steps (events cannot be read back from inside a replayed graph)
    emb.profile = {}
    timed(step_dev, a.steps)
    prof = emb.profile
    emb.profile = None

    use_graph = not a.no_graph
    step_fn, e2e_fn = step_dev, step_e2e
    if use_graph:
        from cloudtik_b200.graph import GraphedTrainStep
        gs = GraphedTrainStep(net, opt, sets[0]["dev"], post_backward=post)

        def step_fn(i):
            return gs.step(*sets[i % len(sets)]["dev"])

        pending = {}

        readers = []

        def e2e_fn(i):
            # every step: H2D of its pinned-host batch (issued one step ahead on a copy stream so that it
            # overlaps the previous step's compute), graph replay, D2H copy of the loss; the host READS each
            # loss one step later (inside the timed region), so the GPU never waits for the host
            k = pending.pop(i) if i in pending else gs.prefetch(sets[i % len(sets)]["host"])
            pending[i + 1] = gs.prefetch(sets[(i + 1) % len(sets)]["host"])
            gs.step_prefetched(k)
            readers.append(gs.loss_to_host_async())
            return readers.pop(0)() if len(readers) > 1 else None

        def e2e_flush():
            while readers:
                readers.pop(0)()
        e2e_fn.flush = e2e_flush

        for i in range(3):
            step_fn(i)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    ms_total, last_loss = timed(step_fn, a.steps)
    loss_val = float(last_loss.cpu())

    for i in range(2):
        e2e_fn(i)
    if use_graph:
        e2e_flush()
        pending.clear()  # the timed region stages its own batches (K+1 H2D copies for K steps)
    ms_e2e, _ = timed(e2e_fn, a.steps)

    # the same end-to-end step fed with RAW records (160 B per sample over PCIe, decoded by b200_criteo_transform)
    ms_rec = None
    if use_graph and world == 1:
        pend = {}

        def rec_fn(i):
            k = pend.pop(i) if i in pend else gs.prefetch_records(sets[i % len(sets)]["host_rec"])
            pend[i + 1] = gs.prefetch_records(sets[(i + 1) % len(sets)]["host_rec"])
            gs.step_records(k)
            readers.append(gs.loss_to_host_async())
            return readers.pop(0)() if len(readers) > 1 else None
        rec_fn.flush = e2e_flush

        for i in range(2):
            rec_fn(i)
        e2e_flush()
        pend.clear()
        ms_rec, _ = timed(rec_fn, a.steps)

    clk = clocks.stop() if clocks else None  # sampled every 20 ms across the three timed legs (all under load)
    if a.timeline:
        write_timeline(a.timeline, rank, step_fn, barrier)
    c5 = None
    if world > 1 and not a.no_c5:
        # BASELINE configs[4] on the same ranks, after the headline legs (their model is released first)
        import gc
        net.emb_l = None
        emb = gs = step_fn = e2e_fn = None
        gc.collect()
        torch.cuda.empty_cache()
        c5 = c5_measure(a, world, rank, dev, min(a.steps, 10), 3)
    if rank != 0:
        return
    ms_step = ms_total / a.steps
    value = B_glob * a.steps / (ms_total / 1e3)
    e2e_value = B_glob * a.steps / (ms_e2e / 1e3)
    h2d = sum(t.numel() * t.element_size() for t in sets[0]["host"])

    # ---- roofline of the dominant hand-written kernel (CUDA events around single launches)
    ew, ea = 4, (2 if bf16 else 4)
    T_loc = len(rows_local)
    nnz = T_loc * B_glob
    U = statistics.mean(sum(s["uniq"]) for s in sets)
    common = nnz * 8 + (T_loc * B_glob + n_chunks) * 8
    alg = {"emb_fwd": U * D * ew + common + T_loc * B_glob * D * ea,
           "emb_bwd_apply": T_loc * B_glob * D * ea + common + 2 * U * D * ew}
    kt = {}
    for k, evs in prof.items():
        if k in alg and evs:  # n_chunks launches per step: their sum is the per-step kernel time
            kt[k] = statistics.mean(x.elapsed_time(y) for x, y in evs) * n_chunks
    peak, pk = peak_hbm()
    roof = None
    if kt:
        dom = max(kt, key=kt.get)
        ach = alg[dom] / (kt[dom] * 1e-3) / 1e9
        roof = {"kernel": dom, "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4), "frac_of_8TBps_nominal": round(ach / 8000.0, 4),
                "peak_source": pk, "traffic": ncu_traffic(dom) if world == 1 else None,
                "traffic_source": "profiles/roofline_traffic.json <- profiles/r2b_ncu_full.summary.md (ncu DRAM read + "
                                  "write bytes per launch of this bench at the FULL 96 GB table sizes, eager launch order)",
                "algorithmic_bytes": int(alg[dom]),
                "kernel_ms": round(kt[dom], 4), "measured_in": "eager pass of the same steps, CUDA events",
                "others": {k: {"ms": round(v, 4), "GBps": round(alg[k] / (v * 1e-3) / 1e9, 1)}
                           for k, v in kt.items() if k != dom}}

    if roof is not None and use_p2p:
        # NVLink side of the same two kernels: (W-1)/W of the pooled rows are WRITTEN to peers by the forward and
        # (W-1)/W of the gradient rows are READ from peers by the update (900 GB/s per direction)
        nv_bytes = T_loc * B_glob * D * ea * (world - 1) / world
        roof["nvlink"] = {"bytes_per_rank_per_direction": int(nv_bytes), "peak": 900.0, "unit": "GB/s", "kernels": {
            k: {"ms": round(v, 4), "GBps": round(nv_bytes / (v * 1e-3) / 1e9, 1), "frac": round(nv_bytes / (v * 1e-3) / 1e9 / 900.0, 3)}
            for k, v in kt.items()}, "note": "rank 0; measured in the eager pass with the side-stream overlap off"}
    bits = max(1, int(sum(rows_local) - 1).bit_length())
    n_relu = (len(LN_BOT) - 1) + (len(LN_TOP) - 2)  # every MLP layer but the sigmoid one: relu_bwd_bias + chunk sum
    # fwd, key tail + keys, radix passes (one-sweep: all-pass histogram + one kernel per pass, up to 2 M keys; else three
    # kernels per pass), dedup (3), update (the cross-block join runs inside it), interaction f/b, MLP backward glue
    # (bf16 path only); with replicated small tables: their forward, index stage and dense-gradient kernel as well
    passes = (bits + 7) // 8
    sort_l = (1 + passes) if nnz <= (2 << 20) else 3 * passes
    launches = 1 + 2 + sort_l + 3 + 1 + 2 + (2 * n_relu if bf16 else 0)
    if rows_repl:
        launches += 1 + 2 + (1 + 2) + 3 + 1  # lookup, key tail + keys, one-sweep sort (<= 16 bits), dedup, dense gradient
    line = {"metric": "dlrm_train_samples_per_sec", "value": round(value, 1), "unit": "samples/s", "n_gpus": world,
            "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": round(ms_step, 4), "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": a.dtype, "data": "synthetic",
            "config": {"workload": "dlrm_criteo_terabyte_c3", "batch_per_gpu": B_loc, "global_batch": B_glob,
                       "micro_batches_per_step": n_chunks,
                       "sparse_dense_boundary": boundary if world > 1 else None,
                       "tables_sharded": len(rows) - len(rows_repl), "tables_replicated": len(rows_repl),
                       "dim": D, "tables": len(rows), "total_rows": int(sum(rows)), "pooling": 1,
                       "table_dtype": "fp32", "mlp": "bot 13-512-256-128 / top 479-1024-1024-512-256-1",
                       "optimizer": "SGD lr 1.0 (fused sparse update + torch SGD on MLPs)",
                       "parallelism": "single" if world == 1 else
                       f"table-wise x{world} ({'fused peer-memory exchange' if use_p2p else 'NCCL all-to-all'}) + data-parallel MLPs",
                       "l2": "inputs larger than L2 (96 GB of tables, uniform random rows)",
                       "rows_cap": a.rows_cap or None,
                       "launch": "whole step replayed as one CUDA graph" if use_graph else "eager"},
            "e2e": {"value": round(e2e_value, 1), "unit": "samples/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e / a.steps, 4)},
            "e2e_records": None if ms_rec is None else {
                "value": round(B_glob * a.steps / (ms_rec / 1e3), 1), "unit": "samples/s",
                "h2d_bytes_per_step": int(sets[0]["host_rec"].numel() * 4), "d2h_bytes_per_step": 4,
                "ms_per_step": round(ms_rec / a.steps, 4),
                "input": "raw int32 Criteo records (CriteoBinDataset format) decoded on the device, one extra launch"},
            "gpu_launches": launches * a.steps, "gpu_launches_per_step": launches,
            "roofline": roof, "clocks": clk, "final_loss": loss_val}
    if parity is not None:
        line["parity"] = parity
    if world == 1:
        # the step's model is not needed any more: drop EVERY name that keeps its 96 GB of tables alive
        net.emb_l = None
        emb = gs = step_fn = e2e_fn = None
    if world == 1 and not a.no_microbench:
        line["embeddingbag_microbench"] = embeddingbag_microbench(net, None)
    if world == 1 and not a.no_c5:
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        c5 = c5_measure(a, world, rank, dev, min(a.steps, 10), 3)
    if c5 is not None:
        line["config5"] = c5
    if world == 1 and not a.no_cpu_baseline:
        r = cpu_reference_run(a.batch, a.cpu_rows_cap, a.cpu_steps, 3, bf16, threads=host_threads())
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    else:
        line["cpu_baseline"] = None
    if world == 1 and bf16 and not a.no_fp32 and not a.global_batch:
        line["fp32"] = fp32_leg(a, rows, sets, dev)
    print(json.dumps(line), flush=True)


def c5_measure(a, world, rank, dev, steps, warmup):
    """BASELINE configs[4]: multi-hot DLRM (pooling ~20), the three largest tables ROW-WISE sharded over the ranks,
    the other 23 table-wise (both exchanges fused into the embedding kernels over peer memory), fused Adagrad on
    every table, bf16 MLPs / interaction / pooled rows.  The whole step is one CUDA graph (the row filter of the
    row-wise path has a fixed capacity).  CUDA-event timed, max over ranks.  On one GPU every table is local.
    Returns the result dict (rank 0) or None."""
    from cloudtik_b200 import ext_dist
    from cloudtik_b200.dlrm import DLRM_Net, enable_peer_exchange, train_step_chunked
    from cloudtik_b200.graph import GraphedTrainStep
    cap = a.rows_cap or (12_000_000 if world == 1 else 0)  # fp32 tables + fp32 Adagrad state: 1 KB per row
    big = sorted(range(len(C3_ROWS)), key=lambda i: -C3_ROWS[i])[:3]
    order = [i for i in range(len(C3_ROWS)) if i not in big] + sorted(big)  # hot (largest) tables last
    rows = [min(C3_ROWS[i], cap) if cap else C3_ROWS[i] for i in order]
    n_rw = 3 if world > 1 else 0
    P, B_loc = a.pooling, a.batch
    B_glob = B_loc * world
    # Adagrad's first steps are +-lr per touched element: keep them small next to U(+-sqrt(1/n)).  MLP rate: with 20
    # rows pooled per bag the tiny tables (3 ... 63 rows, U(+-0.58) entries) give interaction values in the thousands,
    # and at 1e-2 the first top-MLP layer blows up within a few steps whenever the batch composition is unlucky (seen
    # at 4 and 8 GPUs: every prediction saturated, loss 50.45 = 100 x the share of positive targets); the kernels'
    # work does not depend on it
    lr, lr_emb = 1e-4, 1e-3
    np.random.seed(727)
    torch.manual_seed(727 + rank)
    net = DLRM_Net(D, np.array(rows), np.array(LN_BOT), np.array(LN_TOP), sigmoid_top=len(LN_TOP) - 2,
                   np_init_emb_weight=False, device=dev, lr=lr_emb, emb_optimizer="adagrad", bf16=True, n_rowwise=n_rw)
    # non-zero initial accumulator (torch.optim.Adagrad's initial_accumulator_value): from zero, the first steps move
    # every touched element by +-lr whatever the gradient, and with 5 M lookups per table per step the synthetic
    # model saturates within a few steps
    for st in net.emb_l.tables.states + (net.rowwise.emb.tables.states if net.rowwise is not None else []):
        st.fill_(0.1)
    n_tw = len(rows) - n_rw
    local_ids = list(range(n_tw)) if world == 1 else net.local_ln_emb
    if world > 1:
        enable_peer_exchange(net, B_loc)
    dense = [p for p in net.parameters() if p.requires_grad]
    opt = torch.optim.SGD(dense, lr=lr)
    from cloudtik_b200.mlp import enable_bf16_shadows
    enable_bf16_shadows([net.bot_l, net.top_l], opt)
    # Adagrad is not linear in the gradient scale: every rank divides its loss by world, MLP gradients are summed
    sync_dense = ext_dist.DenseGradAllreduce(dense, average=False) if world > 1 else None

    def bags(rs, n_rows_list):
        idx = np.stack([np.floor(rs.rand(B_glob * P) * n).astype(np.int64) for n in n_rows_list])
        off = np.arange(len(n_rows_list) * B_glob + 1, dtype=np.int64) * P
        return idx.reshape(-1), off
    sets = []
    for k in range(2):
        rs_loc = np.random.RandomState(1000 * rank + k)
        rs_glob = np.random.RandomState(99 + k)  # row-wise bags: every rank must see the SAME global batch
        X = rs_loc.rand(B_loc, 13).astype(np.float32)
        Y = np.round(rs_loc.rand(B_loc, 1)).astype(np.float32)
        idx, off = bags(rs_loc, [rows[i] for i in local_ids])
        host = [X, Y, idx, off]
        if n_rw:
            host += list(bags(rs_glob, rows[n_tw:]))
        host = [torch.from_numpy(v).pin_memory() for v in host]
        sets.append({"host": host, "dev": [h.to(dev, non_blocking=True) for h in host]})
    torch.cuda.synchronize()
    scale = 1.0 / world

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        for i in range(n):
            last = fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        return float(ms.item()), last

    use_graph = not a.no_graph
    if use_graph:
        gs = GraphedTrainStep(net, opt, sets[0]["dev"], post_backward=sync_dense, loss_scale=scale)

        def step(i):
            return gs.step(*sets[i % 2]["dev"])
    else:
        def step(i):
            return train_step_chunked(net, opt, sets[i % 2]["dev"], sync_dense, scale)
    for i in range(max(warmup, 3)):
        step(i)
    ms_total, last = timed(step, steps)
    loss_val = float(last.cpu())
    # end to end: pinned-host inputs copied in every step, loss read back every step
    stage = [torch.empty_like(t) for t in sets[0]["dev"]]

    def step_e2e(i):
        for d, h in zip(stage, sets[i % 2]["host"]):
            d.copy_(h, non_blocking=True)
        E = gs.step(*stage) if use_graph else train_step_chunked(net, opt, stage, sync_dense, scale)
        return float(E.cpu())
    step_e2e(0)
    ms_e2e, _ = timed(step_e2e, steps)
    # kernel-level roofline of the multi-hot update (the dominant embedding kernel of this config): an eager pass
    # with CUDA events around the launches (local tables of this rank: 23/W table-wise ones)
    emb = net.emb_l
    emb.profile = {}
    timed(lambda i: train_step_chunked(net, opt, sets[i % 2]["dev"], sync_dense, scale), min(steps, 5))
    prof, emb.profile = emb.profile, None
    overflow = int(net.rowwise.error_flag.item()) if (net.rowwise is not None and net.rowwise.error_flag is not None) else 0
    if rank != 0:
        return None
    T_loc = len(local_ids)
    nnz = T_loc * B_glob * P
    roof = None
    if prof.get("emb_bwd_apply"):
        kms = statistics.mean(x.elapsed_time(y) for x, y in prof["emb_bwd_apply"])
        # distinct rows touched per step, per table: n * (1 - exp(-lookups / n)) for uniform ids
        U = sum(rows[i] * (1.0 - np.exp(-B_glob * P / rows[i])) for i in local_ids)
        # bf16 gradient rows + ids + offsets + (fp32 row + fp32 Adagrad state) read and written
        alg = T_loc * B_glob * D * 2 + nnz * 8 + (T_loc * B_glob + 1) * 8 + 2 * U * D * 4 + 2 * U * D * 4
        peak, pk = peak_hbm()
        roof = {"kernel": "emb_bwd_apply (multi-hot, fused Adagrad, table-wise tables of rank 0)", "bound": "hbm",
                "achieved": round(alg / (kms * 1e-3) / 1e9, 1), "peak": peak, "unit": "GB/s",
                "frac": round(alg / (kms * 1e-3) / 1e9 / peak, 4), "peak_source": pk, "kernel_ms": round(kms, 4),
                "algorithmic_bytes": int(alg), "distinct_rows_estimate": int(U),
                "others": {k: round(statistics.mean(x.elapsed_time(y) for x, y in v), 4)
                           for k, v in prof.items() if k != "emb_bwd_apply" and v}}
    h2d = sum(t.numel() * t.element_size() for t in sets[0]["host"])
    return {"metric": "dlrm_train_samples_per_sec", "value": round(B_glob * steps / (ms_total / 1e3), 1),
            "unit": "samples/s", "n_gpus": world, "steps": steps, "warmup": max(warmup, 3),
            "ms_per_step": round(ms_total / steps, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "dlrm_multihot_rowwise_adagrad_c5", "batch_per_gpu": B_loc, "global_batch": B_glob,
                       "dim": D, "tables": len(rows), "pooling": P, "rows_cap": cap or None,
                       "lookups_per_step": int(len(rows) * B_glob * P),
                       "row_wise_tables": n_rw, "row_wise_rows": [int(r) for r in rows[n_tw:]] if n_rw else [],
                       "optimizer": "fused sparse Adagrad (tables, initial accumulator 0.1) + torch SGD (MLPs)",
                       "table_dtype": "fp32",
                       "parallelism": "single" if world == 1 else
                       f"table-wise x{world} + row-wise x{world} (both exchanges fused over peer memory) + "
                       "data-parallel MLPs",
                       "l2": "inputs larger than L2",
                       "launch": "whole step replayed as one CUDA graph" if use_graph else "eager",
                       "row_filter_overflow": overflow},
            "e2e": {"value": round(B_glob * steps / (ms_e2e / 1e3), 1), "unit": "samples/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e / steps, 4)},
            "final_loss": loss_val, "roofline": roof}


def run_c5_arm(a):
    from cloudtik_b200 import ext_dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        ext_dist.init_distributed(rank=rank, size=world, backend="nccl")
    clocks = ClockSampler(local_rank) if rank == 0 else None
    line = c5_measure(a, world, rank, dev, a.steps, a.warmup)
    clk = clocks.stop() if clocks else None
    if rank == 0:
        line["clocks"] = clk
        line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)


def main():
    a = parse()
    if a.impl == "reference":
        run_reference_arm(a)
    elif a.workload == "c5":
        run_c5_arm(a)
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.destroy_process_group()
    else:
        run_b200_arm(a)
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
