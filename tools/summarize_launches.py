"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.

    python tools/summarize_launches.py profiles/<file>.csv [--steps K] > profiles/<file>.summary.md
"""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 1
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    total = 0.0
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1000 if u == "ns" else v * 1000 if u == "ms" else v * 1e6 if u == "s" else v
        name = row["Kernel Name"].split("(")[0][:100]
        c = agg.setdefault(name, [0, 0.0])
        c[0] += 1
        c[1] += v
        total += v
    print(f"# {path}: {sum(c[0] for c in agg.values())} launches, {total / 1000:.3f} ms total "
          f"(cold-cache, serialised under ncu; compare shares, not absolutes)\n")
    print("| kernel | launches | total us | us/launch | share |")
    print("|---|---:|---:|---:|---:|")
    for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {n} | {t:.1f} | {t / n:.1f} | {100 * t / total:.1f}% |")
    ours = sum(t for k, (n, t) in agg.items() if "b200::" in k)
    tail = (f" ({steps} step(s) captured => {total / 1000 / steps:.3f} ms/step serialised)" if "--steps" in sys.argv
            else " (the capture window is a launch count, not a whole number of steps)")
    print(f"\nb200:: kernels: {ours / 1000:.3f} ms = {100 * ours / total:.1f}% of the captured time" + tail)


if __name__ == "__main__":
    main()
