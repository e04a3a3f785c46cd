#!/usr/bin/env python3
"""Where do the persistent engine's CTAs wait?  Runs one grid under several env variants and prints
device-timed GLUPS plus the per-CTA cycle split (dev tool, GPU box).
    python tools/probe_sync.py --synthetic 16384x256 --ngpus 2 --variants "LBM_EDGE_WEIGHT_16=16;LBM_EDGE_WEIGHT_16=24"
"""
import argparse
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
lbm = importlib.import_module("lattice-boltzmann_b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--synthetic", default="16384x256")
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--ngpus", type=int, default=1)
    ap.add_argument("--variants", default="LBM_ENGINE=0", help="semicolon-separated variants, each a comma-separated list of ENV=value")
    ap.add_argument("--profile", type=int, default=1)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--dump", type=int, default=0)
    ap.add_argument("--first-device", type=int, default=0)
    ap.add_argument("--ctas-per-sm", type=int, default=0)
    args = ap.parse_args()
    nx, ny = (int(v) for v in args.synthetic.split("x"))
    p = dict(nx=nx, ny=ny, density=0.1, accel=0.01, omega=1.85)
    mask = np.zeros((ny, nx), np.int32)
    mask[0, :] = mask[-1, :] = 1
    mask[:, 0] = mask[:, -1] = 1
    mask[:, nx // 4] = mask[:, nx // 2] = 1
    lines = int(mask.sum())
    cells = nx * ny
    for var in args.variants.split(";"):
        saved = {}
        for kv in var.split(","):
            if not kv:
                continue
            k, v = kv.split("=")
            saved[k] = os.environ.get(k)
            os.environ[k] = v
        os.environ["LBM_PROFILE"] = str(args.profile)
        try:
            lat = lbm.Lattice(nx, ny, args.steps * (args.reps + 1), p["density"], p["accel"], p["omega"], mask,
                              obstacle_lines=lines, ngpus=args.ngpus, threads_per_cta=args.threads, ctas_per_sm=args.ctas_per_sm, first_device=args.first_device)
            lat.run(args.steps); lat.sync()
            best = 1e30
            for _ in range(args.reps):
                lat.run(args.steps); lat.sync()
                best = min(best, lat.last_run_ms())
            glups = cells * args.steps / (best * 1e-3) / 1e9
            print(f"{args.synthetic} x{args.ngpus} [{var}] [{lat.describe()}] {best / args.steps * 1e3:.2f} us/step "
                  f"{glups:.2f} GLUPS ({glups / args.ngpus:.2f}/GPU)", flush=True)
            if args.profile:
              for slab in range(args.ngpus):
                  pr = lat.profile(slab).astype(np.float64)        # [cta, warp, (wait, barrier, halo pull, total)]
                  tot = pr[:, :, 3].mean()
                  wait = pr[:, :, 0].mean(axis=1) / tot * 100    # per CTA, % of run
                  bar = pr[:, :, 1].mean(axis=1) / tot * 100
                  ll = pr[:, :, 2].mean(axis=1) / tot * 100
                  busy = 100 - wait - ll - bar
                  edge = ll > 0
                  def st(a):
                      return f"mean {a.mean():5.2f} p50 {np.median(a):5.2f} max {a.max():5.2f}"
                  print(f"   cycles/step {tot / args.steps:.0f};  % of run per CTA: wait[{st(wait)}]  barrier[{st(bar)}]  busy[{st(busy)}]")
                  if edge.any() and not edge.all():
                      print(f"   {int(edge.sum())} edge CTAs: wait[{st(wait[edge])}] barrier[{st(bar[edge])}] halo-pull[{st(ll[edge])}] busy[{st(busy[edge])}]")
                      print(f"   others: wait[{st(wait[~edge])}] barrier[{st(bar[~edge])}] busy[{st(busy[~edge])}]")
                  w0 = pr[:, 0, 1].mean() / tot * 100
                  wr = pr[:, 1:, 1].mean() / tot * 100
                  print(f"   barrier wait warp0 {w0:.2f}%  other warps {wr:.2f}%")
                  if args.dump:
                      for i in range(0, min(len(busy), args.dump)):
                          print(f"      cta {i:3d}{'E' if edge[i] else ' '} wait {wait[i]:5.1f} sig {bar[i]:4.1f} ll {ll[i]:5.1f} busy {busy[i]:5.1f} | per-warp wait% " +
                                " ".join(f"{v:4.0f}" for v in pr[i, :, 0] / tot * 100))
                  top = np.argsort(-busy)[:8]
                  print("   busiest CTAs:", ", ".join(f"{i}:{busy[i]:.1f}%{'E' if edge[i] else ''}" for i in top))
            lat.close()
        except lbm.LbmError as e:
            print(var, "ERROR", e)
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


if __name__ == "__main__":
    main()
