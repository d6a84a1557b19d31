"""Condense an `ncu --set full` report into the few numbers the roofline discussion needs.

    ncu -i gpurun_out/r1f_kernels.ncu-rep --page raw --csv > /tmp/raw.csv
    python tools/ncu_summary.py /tmp/raw.csv profiles/r1f_ncu_full.summary.md [profiles/roofline_traffic.json]

One row per captured launch: duration, DRAM bytes read / written, DRAM throughput, L2 hit rate, achieved
occupancy, registers, grid, tensor-pipe activity (tcgen05 kernels).  With a third argument the DRAM bytes per
launch are also written as JSON under the names bench.py uses for its `roofline.traffic` field
(emb_fwd = emb_fwd_kernel, emb_bwd_apply = window_update_kernel + window_join_kernel of one step).
"""
import csv
import json
import sys

COLS = [
    ("gpu__time_duration.sum", "us"),
    ("dram__bytes_read.sum", "DRAM rd MB"),
    ("dram__bytes_write.sum", "DRAM wr MB"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe % (active)"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe % (elapsed)"),
    ("smsp__sass_inst_executed_op_tmem_ldt.sum", "tcgen05.ld insts"),
]


def to_mb(v, unit):
    f = float(v)
    return f * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(unit, 1.0)


def main():
    raw, out_md = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(open(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    lines = ["| kernel | " + " | ".join(n for _, n in COLS) + " | DRAM TB/s |", "|---|" + "---:|" * (len(COLS) + 1)]
    traffic = {}
    for r in data:
        name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").strip()
        cells = []
        rd = wr = 0.0
        for key, label in COLS:
            if key not in ix or r[ix[key]] in ("", "n/a"):
                cells.append("-")
                continue
            v, u = r[ix[key]], units[ix[key]]
            if "bytes" in key:
                mb = to_mb(v, u)
                cells.append(f"{mb:.1f}")
                if "read" in key:
                    rd = mb
                else:
                    wr = mb
            else:
                f = float(v)
                cells.append(f"{f:.0f}" if f == int(f) else f"{f:.2f}")
        us = float(r[ix["gpu__time_duration.sum"]])
        lines.append(f"| `{name}` | " + " | ".join(cells) + f" | {(rd + wr) / us:.2f} |")
        key = "emb_fwd" if "emb_fwd_kernel" in name else ("emb_bwd_apply" if "window_" in name else name)
        traffic[key] = traffic.get(key, 0) + int((rd + wr) * 1e6)
    open(out_md, "w").write("\n".join(lines) + "\n")
    if len(sys.argv) > 3:
        json.dump(traffic, open(sys.argv[3], "w"), indent=1, sort_keys=True)
        open(sys.argv[3], "a").write("\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
