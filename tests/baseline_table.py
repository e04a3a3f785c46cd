#!/usr/bin/env python
"""Fills BASELINE.md section 4's table on the GPU box: for every BASELINE.json configuration that fits
ONE GPU it runs the drop-in program (./d2q9-bgk, full length) and judges its files with tools/check.py
against the reference's shipped known answers (tests/golden/shipped_*.npz); beside it the unmodified
reference's OpenMP build (oracle/_ref/ref-omp) on the box's host cores (full length up to 1024x1024, a
bounded sample above) and, for 128x128, the serial build (oracle/_ref/ref-serial), judged the same way.
Writes gpurun_out/baseline_table.md.  Test infrastructure (lives under tests/): besides bench.py the only script that runs
oracle/_ref binaries, as the measured CPU baseline only.

usage: python tests/baseline_table.py [--skip-large]
"""
import argparse
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench          # noqa: E402  (helpers: grid_files, host_cores, run_reference_sample)
import oracle_lib     # noqa: E402  (format_final_state / format_av_vels)

EXE = os.path.join(ROOT, "d2q9-bgk")
REF_SERIAL = os.path.join(ROOT, "oracle", "_ref", "ref-serial")
CHECK = os.path.join(ROOT, "tools", "check.py")
BYTES = 72.0


def compute_seconds(stdout):
    return float(re.search(r"Elapsed Compute time:\s+([0-9.]+)", stdout).group(1))


def judge(tmp, grid):
    """tools/check.py (the reference checker's rule: max 1 % difference) against the shipped known answers."""
    path = os.path.join(ROOT, "tests", "golden", f"shipped_{grid}.npz")
    if not os.path.exists(path):
        return "no shipped answers"
    g = np.load(path)
    nx, ny = (int(v) for v in grid.split("x"))
    with open(os.path.join(tmp, "ref_av.dat"), "w") as fh:
        fh.write(oracle_lib.format_av_vels(g["av_vels"]))
    args = [sys.executable, CHECK, "--ref-av-vels-file", os.path.join(tmp, "ref_av.dat"),
            "--av-vels-file", os.path.join(tmp, "av_vels.dat")]
    what = "av_vels"
    if "pressure" in g.files:
        # obstacle column of the reference file: taken from the produced file (the checker compares pressure only)
        obst = np.loadtxt(os.path.join(tmp, "final_state.dat"), usecols=6, dtype=np.int32).reshape(ny, nx)
        z = np.zeros((ny, nx), np.float32)
        with open(os.path.join(tmp, "ref_fs.dat"), "w") as fh:
            fh.write(oracle_lib.format_final_state(z, z, z, g["pressure"], obst))
        args += ["--ref-final-state-file", os.path.join(tmp, "ref_fs.dat"), "--final-state-file",
                 os.path.join(tmp, "final_state.dat")]
        what = "av_vels + final_state"
    else:
        args += ["--ref-final-state-file", os.path.join(tmp, "final_state.dat"), "--final-state-file",
                 os.path.join(tmp, "final_state.dat")]
    r = subprocess.run(args, capture_output=True, text=True)
    ok = r.returncode == 0 and "Both tests passed!" in r.stdout
    return ("PASS" if ok else "FAIL") + f" ({what} vs shipped)"


def run_program(exe, grid, env=None):
    pfile, ofile = bench.grid_files(grid)
    with tempfile.TemporaryDirectory() as tmp:
        out = subprocess.run([exe, pfile, ofile], cwd=tmp, env=dict(os.environ, **(env or {})), check=True,
                             capture_output=True, text=True).stdout
        secs = compute_seconds(out)
        nx, ny = (int(v) for v in grid.split("x"))
        small = nx * ny <= 1024 * 1024
        verdict = judge(tmp, grid) if small else "-"
    return secs, verdict


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-large", action="store_true")
    args = ap.parse_args()
    cores = bench.host_cores()
    peak = bench.hbm_peak()[0]
    rows = []

    def iters_of(grid):
        with open(bench.grid_files(grid)[0]) as fh:
            return int(fh.read().split()[2])

    def glups(grid, secs, iters=None):
        nx, ny = (int(v) for v in grid.split("x"))
        return nx * ny * (iters or iters_of(grid)) / secs / 1e9

    # config 1: the reference's serial build on one host core, 128x128, judged by the checker
    s, v = run_program(REF_SERIAL, "128x128")
    rows.append(("128x128", "reference, serial, 1 host core", f"{glups('128x128', s):.4f} ({s:.2f} s)", "-", "-", v))
    print(rows[-1], flush=True)

    grids = ["128x128", "128x256", "256x256", "1024x1024"] + ([] if args.skip_large else ["16384x1024"])
    for grid in grids:
        s, v = run_program(EXE, grid, {"LBM_NGPUS": "1"})
        g = glups(grid, s)
        nx, ny = (int(x) for x in grid.split("x"))
        if nx * ny <= 1024 * 1024:
            cs, cg = bench.run_reference_sample(grid, iters_of(grid), cores)
            cpu = f"{cg:.3f} ({cores} cores, full length, {cs:.1f} s)"
        else:
            it = bench.reference_sample_iters(grid, cores, 8.0)
            cs, cg = bench.run_reference_sample(grid, it, cores)
            cpu = f"{cg:.3f} ({cores} cores, maxIters={it}, {cs:.1f} s)"
        rows.append((grid, "1 GPU, ./d2q9-bgk full length", f"{g:.2f} ({s:.3f} s)", f"{100 * g * BYTES / peak:.0f} %", cpu, v))
        print(rows[-1], flush=True)

    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "baseline_table.md"), "w") as fh:
        fh.write("| Config | What ran | GLUPS (compute time) | % of measured HBM roofline | reference OpenMP GLUPS | check.py |\n")
        fh.write("|---|---|---|---|---|---|\n")
        for r in rows:
            fh.write("| " + " | ".join(r) + " |\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
