#!/usr/bin/env python3
"""Summarises ncu captures (run here, no GPU needed) into markdown for profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches_bench.csv
  python tools/ncu_summary.py full gpurun_out/prof_persistent_16384x1024.ncu-rep <steps> <cells>
"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor", "sm__cycles_elapsed.avg",
        "smsp__cycles_active.avg", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.sum", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum"]


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
    g, b = hdr.index("Grid Size"), hdr.index("Block Size")
    agg = defaultdict(lambda: [0, 0.0, "", ""])
    for r in rows[1:]:
        name = r[k].split("(")[0]
        a = agg[name]
        a[0] += 1
        a[1] += float(r[v].replace(",", ""))
        a[2], a[3] = r[g], r[b]
    tot = sum(a[1] for a in agg.values())
    print("| kernel | launches | grid | block | total ms | share |\n|---|---|---|---|---|---|")
    for name, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| `{name}` | {a[0]} | {a[2]} | {a[3]} | {a[1] / 1e6:.3f} | {100 * a[1] / tot:.2f} % |")
    print(f"\ntotal device time in kernels: {tot / 1e6:.3f} ms over {sum(a[0] for a in agg.values())} launches")


def full(path, steps, cells):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {h: (vals[i], units[i]) for i, h in enumerate(hdr)}
    print(f"kernel: `{d['Kernel Name'][0].split('(')[0]}`  grid {d['Grid Size'][0]} block {d['Block Size'][0]}\n")
    print("| metric | value | unit |\n|---|---|---|")
    for key in KEEP:
        if key in d:
            print(f"| {key} | {d[key][0]} | {d[key][1]} |")
    stalls = {h: float(v[0]) for h, v in d.items() if h.startswith("smsp__average_warps_issue_stalled") and v[0] not in ("", "n/a")}
    print("\nwarp stall reasons (warps stalled per issue-active cycle), top 8:\n")
    for h, v in sorted(stalls.items(), key=lambda x: -x[1])[:8]:
        print(f"* {h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')}: {v:.3f}")
    def num(key):
        v, u = d[key]
        f = float(v.replace(",", ""))
        return f * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0, "second": 1.0,
                    "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9}.get(u, 1)
    t = num("gpu__time_duration.sum")
    dram = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
    alg = 72.0 * steps * cells
    inst = float(d["smsp__inst_executed.sum"][0].replace(",", ""))
    print(f"\nderived: launch = {steps} timesteps x {cells} cells; duration under ncu {t * 1e3:.3f} ms "
          f"({t / steps * 1e6:.2f} us/step, cold caches, serialised)")
    print(f"* algorithmic bytes per launch = 72 x {steps} x {cells} = {alg / 1e9:.3f} GB -> {alg / t / 1e9:.0f} GB/s under ncu")
    print(f"* DRAM traffic per launch (dram__bytes_read+write) = {dram / 1e9:.3f} GB = {dram / alg:.3f} x algorithmic "
          f"({dram / steps / cells:.1f} B per cell update)")
    print(f"* warp instructions per 128-cell task = {inst / (steps * cells / 128):.0f}  ({inst / (steps * cells) * 32 / 32:.2f} per cell per lane-quad x4)")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        full(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]))
