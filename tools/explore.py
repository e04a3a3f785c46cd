#!/usr/bin/env python3
"""Sweeps engine configurations on one grid and prints device-timed GLUPS (dev tool, GPU box)."""
import argparse
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
lbm = importlib.import_module("lattice-boltzmann_b200")


def hbm_gbs():
    import json
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"])
    except Exception:
        return 6650.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", default="1024x1024")
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--ngpus", type=int, default=1)
    ap.add_argument("--synthetic", default=None, help="NXxNY: closed box + walls at nx/4, nx/2 instead of a shipped grid")
    ap.add_argument("--configs", default="1:256:2:1,1:512:1:1,1:256:1:1,1:128:4:1,1:256:2:2,2:256:0:1")
    args = ap.parse_args()
    pfile = os.path.join(ROOT, "inputs", f"input_{args.grid}.params")
    ofile = os.path.join(ROOT, "inputs", f"obstacles_{args.grid}.dat")
    if args.synthetic:
        import numpy as np
        nx, ny = (int(v) for v in args.synthetic.split("x"))
        p = dict(nx=nx, ny=ny, density=0.1, accel=0.01, omega=1.85)
        mask = np.zeros((ny, nx), np.int32)
        mask[0, :] = mask[-1, :] = 1
        mask[:, 0] = mask[:, -1] = 1
        mask[:, nx // 4] = mask[:, nx // 2] = 1
        lines = int(mask.sum())
        args.grid = args.synthetic
    else:
        p = lbm.read_params(pfile)
        mask, lines = lbm.read_obstacles(ofile, p["nx"], p["ny"])
    cells = p["nx"] * p["ny"]
    for cfg in args.configs.split(","):
        engine, threads, cps, fp = (int(v) for v in cfg.split(":"))
        try:
            lat = lbm.Lattice(p["nx"], p["ny"], args.steps * (args.reps + 1), p["density"], p["accel"], p["omega"], mask,
                              obstacle_lines=lines, ngpus=args.ngpus, engine=engine, threads_per_cta=threads,
                              ctas_per_sm=cps, strict_fp=fp)
        except lbm.LbmError as e:
            print(cfg, "ERROR", e)
            continue
        lat.run(args.steps); lat.sync()
        best = 1e30
        for _ in range(args.reps):
            lat.run(args.steps); lat.sync()
            best = min(best, lat.last_run_ms())
        glups = cells * args.steps / (best * 1e-3) / 1e9
        print(f"{args.grid} cfg={cfg} [{lat.describe()}] {best / args.steps * 1e3:.2f} us/step  {glups:.2f} GLUPS  "
              f"{glups * 72 / hbm_gbs() * 100 / args.ngpus:.1f}% of measured HBM roofline", flush=True)
        lat.close()


if __name__ == "__main__":
    main()
