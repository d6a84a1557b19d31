"""Per-rank step timeline from the files `bench.py --timeline <path>` writes (<path>.rank<r>.json).

    python tools/timeline_summary.py gpurun_out/tl_n8 [--md]

For every rank: the span of one steady-state step, the busy time of each stream, the time spent inside
cross-rank barrier kernels (= waiting for a slower rank or for data in flight), the NCCL time, and the longest
kernels by stream.  The LAST rank to reach a barrier is the one with the shortest barrier time.
"""
import glob
import json
import sys
from collections import defaultdict


def load(prefix):
    out = []
    for f in sorted(glob.glob(prefix + ".rank*.json")):
        out.append(json.load(open(f)))
    return sorted(out, key=lambda d: d["rank"])


def short(name):
    for pat, s in (("symm_mem", "barrier"), ("barrier", "barrier"), ("nccl", "nccl"), ("ncclDevKernel", "nccl")):
        if pat in name:
            return s
    n = name.replace("void ", "").replace("b200::", "b200::")
    return n.split("<")[0].split("(")[0][:48]


def steps_of(events, nsteps):
    """Cut the event list into steps at the first kernel of each graph replay: the step starts with the same
    kernel name every time (the first event); use its occurrences."""
    first = events[0]["name"]
    starts = [i for i, e in enumerate(events) if e["name"] == first]
    # the first kernel may occur several times per step: keep every len(starts)/nsteps-th
    per = max(1, len(starts) // nsteps)
    starts = starts[::per][:nsteps]
    cuts = starts + [len(events)]
    return [events[cuts[i]:cuts[i + 1]] for i in range(len(starts))]


def summarise(d):
    ev = [e for e in d["events"] if e["dur_us"] is not None]
    st = steps_of(ev, d["steps"])
    step = st[len(st) // 2]  # a middle step
    t0 = step[0]["ts_us"]
    t1 = max(e["ts_us"] + e["dur_us"] for e in step)
    by_stream = defaultdict(float)
    by_kind = defaultdict(float)
    for e in step:
        by_stream[e["stream"]] += e["dur_us"]
        by_kind[short(e["name"])] += e["dur_us"]
    return {"rank": d["rank"], "span_us": t1 - t0, "kernels": len(step), "streams": dict(by_stream),
            "barrier_us": by_kind.get("barrier", 0.0), "nccl_us": by_kind.get("nccl", 0.0),
            "top": sorted(by_kind.items(), key=lambda kv: -kv[1])[:12], "step": step, "t0": t0}


def main():
    prefix = sys.argv[1]
    ranks = load(prefix)
    if not ranks:
        sys.exit("no timeline files for " + prefix)
    rows = [summarise(d) for d in ranks]
    print("| rank | step span us | kernels | barrier-kernel us (wait) | nccl us | busy us per stream |")
    print("|---:|---:|---:|---:|---:|---|")
    for r in rows:
        streams = ", ".join(f"s{k}: {v:.0f}" for k, v in sorted(r["streams"].items(), key=lambda kv: -kv[1]))
        print(f"| {r['rank']} | {r['span_us']:.0f} | {r['kernels']} | {r['barrier_us']:.0f} | {r['nccl_us']:.0f} | {streams} |")
    last = min(rows, key=lambda r: r["barrier_us"])
    print(f"\nlast rank (least time waiting in barriers): {last['rank']}")
    for r in rows:
        if r["rank"] in (0, last["rank"]) or "--all" in sys.argv:
            print(f"\n### rank {r['rank']}: kernels of one step (start us relative to the step, duration us, stream)\n")
            for e in r["step"]:
                if e["dur_us"] >= 8 or "--all" in sys.argv:
                    print(f"{e['ts_us'] - r['t0']:9.1f} {e['dur_us']:8.1f}  s{e['stream']}  {short(e['name'])}")


if __name__ == "__main__":
    main()
