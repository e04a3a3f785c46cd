#!/usr/bin/env python3
"""Turns the ncu artefacts of tools/profile_r01.sh / profile_r02.sh (gpurun_out/) into the tracked summaries under profiles/.
    python tools/ncu_summary.py r01        # runs here (no GPU needed): ncu -i <rep> --page raw --csv
"""
import csv
import io
import json
import os
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor", "sm__cycles_elapsed.avg",
        "smsp__cycles_active.avg", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "launch__shared_mem_per_block_dynamic"]


def raw_page(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    header, units, vals = rows[0], rows[1], rows[2]
    return OrderedDict((h, (v, u)) for h, u, v in zip(header, units, vals))


def to_bytes(v, u):
    f = float(v.replace(",", ""))
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[u]


def to_ms(v, u):
    f = float(v.replace(",", ""))
    return f * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}[u]


def full_summary(tag, grid, steps, kind="persistent", config="1:256:2:1"):
    rep = os.path.join(OUT, f"prof_{kind}_{grid}.ncu-rep")
    m = raw_page(rep)
    nx, ny = (int(v) for v in grid.split("x"))
    cells = nx * ny
    lines = [f"# {tag} — ncu --set full of lbm_{kind}_kernel<strict> on {grid} ({steps} timesteps in one launch)", "",
             f"Command: `python tools/explore.py --grid {grid} --steps {steps} --reps 1 --configs {config}` (2nd launch captured, "
             f"`-s 1 -c 1`), ncu --set full --clock-control none --import-source on (tools/profile_{tag}.sh). "
             f"Report: gpurun_out/prof_{kind}_{grid}.ncu-rep (scratch).", "",
             f"kernel: `{m['Kernel Name'][0]}`  grid {m.get('Grid Size', ('?',))[0]} block {m.get('Block Size', ('?',))[0]}", "",
             "| metric | value | unit |", "|---|---|---|"]
    for k in WANT:
        if k in m:
            lines.append(f"| {k} | {m[k][0]} | {m[k][1]} |")
    stalls = sorted(((float(v[0].replace(",", "")), k.split("issue_stalled_")[1].split("_per_warp")[0]) for k, v in m.items()
                     if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")),
                    reverse=True)[:8]
    if not stalls:
        stalls = sorted(((float(v[0].replace(",", "")), k.replace("smsp__average_warp_latency_issue_stalled_", "").replace(".ratio", ""))
                         for k, v in m.items() if k.startswith("smsp__average_warp_latency_issue_stalled_")), reverse=True)[:8]
    lines += ["", "warp stall reasons (warps stalled per issue-active cycle), top 8:", ""]
    lines += [f"* {n}: {v:.3f}" for v, n in stalls]
    ms = to_ms(*m["gpu__time_duration.sum"])
    rd, wr = to_bytes(*m["dram__bytes_read.sum"]), to_bytes(*m["dram__bytes_write.sum"])
    alg = 72.0 * steps * cells
    inst = float(m["smsp__inst_executed.sum"][0].replace(",", ""))
    tasks = steps * ny * ((nx + 127) // 128)
    lines += ["", f"derived: launch = {steps} timesteps x {cells} cells; duration under ncu {ms:.3f} ms "
                  f"({ms / steps * 1e3:.2f} us/step, cold caches, serialised)",
              f"* algorithmic bytes per launch = 72 x {steps} x {cells} = {alg / 1e9:.3f} GB -> {alg / ms / 1e6:.0f} GB/s under ncu",
              f"* DRAM traffic per launch (dram__bytes_read+write) = {(rd + wr) / 1e9:.3f} GB = {(rd + wr) / alg:.3f} x algorithmic "
              f"({(rd + wr) / (steps * cells):.1f} B per cell update)",
              f"* warp instructions per 128-cell task = {inst / tasks:.0f}  ({inst / tasks / 128:.2f} per cell per lane-quad x4)"]
    with open(os.path.join(PROF, f"{tag}_ncu_{kind}_{grid}.md"), "w") as fh:
        fh.write("\n".join(lines) + "\n")
    return (rd + wr) / (steps * cells)


def launch_list(tag):
    src = os.path.join(OUT, "launches_bench.csv")
    rows = []
    with open(src) as fh:
        txt = [l for l in fh if not l.startswith("==")]
    for r in csv.DictReader(io.StringIO("".join(txt))):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            ms = to_ms(r["Metric Value"], r["Metric Unit"])
            rows.append((r["Kernel Name"], r["Grid Size"], r["Block Size"], ms))
    agg = OrderedDict()
    for name, g, b, ms in rows:
        key = (name, g, b)
        n, t = agg.get(key, (0, 0.0))
        agg[key] = (n + 1, t + ms)
    tot = sum(t for _, t in agg.values())
    cmd = "python bench.py --steps 1 --warmup 3 --iters 300 --no-cpu-baseline" + ("" if tag == "r01" else " --no-also")
    lines = [f"# {tag} — launch list of `{cmd}`", "",
             f"ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES). "
             f"Raw CSV: {tag}_launches_bench.csv", "", "| kernel | launches | grid | block | total ms | share |", "|---|---|---|---|---|---|"]
    for (name, g, b), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"| `{name[:110]}` | {n} | {g} | {b} | {t:.3f} | {100 * t / tot:.2f} % |")
    lines += ["", f"total device time in kernels: {tot:.3f} ms over {len(rows)} launches", "",
              "The timestep kernel (one persistent launch per lbm_run) is the only kernel that matters in a bench run;",
              "k_fill_uniform = initial state, k_fold_partials = av_vels fold, the at:: kernels are bench.py's L2 flush."]
    with open(os.path.join(PROF, f"{tag}_launches.md"), "w") as fh:
        fh.write("\n".join(lines) + "\n")
    with open(os.path.join(PROF, f"{tag}_launches_bench.csv"), "w") as fh:
        fh.write("".join(txt))


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
    if len(sys.argv) > 2 and sys.argv[2] == "row":       # engine 4 captures of tools/profile_r02_row.sh only
        for grid in ("128x128", "256x256"):
            full_summary(tag, grid, 2000, kind="row", config="4:0:0:1")
        return
    launch_list(tag)
    per = {}
    for grid, steps in (("16384x1024", 40), ("1024x1024", 300)):
        per[grid] = round(full_summary(tag, grid, steps), 2)
    if os.path.exists(os.path.join(OUT, "prof_dataflow_256x256.ncu-rep")):      # engine 3, L2-resident: not in traffic.json
        full_summary(tag, "256x256", 2000, kind="dataflow", config="3:32:0:1")
    with open(os.path.join(PROF, "traffic.json"), "w") as fh:
        json.dump({"source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one lbm_persistent_kernel launch / "
                             f"(timesteps x cells) - see profiles/{tag}_ncu_persistent_<grid>.md",
                   "dram_bytes_per_cell_update": per}, fh, indent=1)
        fh.write("\n")
    print(per)


if __name__ == "__main__":
    main()
