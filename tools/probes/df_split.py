#!/usr/bin/env python3
"""Where does a timestep of the dataflow engine (engine 3) go?  Needs a library built with -DLBM_DF_PROBE
(nvcc ... -DLBM_DF_PROBE -o scratch/probe.so) selected through LBM_LIBRARY, and LBM_PROFILE=1.  Per task and step:
cycles from the first record load to the successful vote (poll), extra polls, cycles from the vote to the last
record store (collide + store issue)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ["LBM_PROFILE"] = "1"
lbm = importlib.import_module("lattice-boltzmann_b200")
for grid in sys.argv[1:] or ["128x128", "256x256", "512x512"]:
    nx, ny = (int(v) for v in grid.split("x"))
    mask = np.zeros((ny, nx), np.int32); mask[0, :] = mask[-1, :] = 1; mask[:, 0] = mask[:, -1] = 1
    steps = 4000
    if nx <= 256:                                              # engine 4: [row warp, (poll, extra polls, barrier wait, collide + stores)]
        with lbm.Lattice(nx, ny, 2 * steps, 0.1, 0.005, 1.85, mask, engine=4) as lat:
            lat.run(steps); lat.sync()
            lat.run(steps); lat.sync()
            ms = lat.last_run_ms()
            pr = lat.profile(0)[:, 0, :].astype(np.float64) / steps
        print(f"{grid} engine 4: {ms / steps * 1e3:.2f} us/step = {ms / steps * 1e-3 * 1.965e9:.0f} cycles | per warp and step: poll {pr[:, 0].mean():.0f} "
              f"(min {pr[:, 0].min():.0f}, max {pr[:, 0].max():.0f}), extra polls {pr[:, 1].mean():.2f}, barrier wait {pr[:, 2].mean():.0f}, "
              f"collide + record stores {pr[:, 3].mean():.0f} (min {pr[:, 3].min():.0f}, max {pr[:, 3].max():.0f})", flush=True)
    with lbm.Lattice(nx, ny, 2 * steps, 0.1, 0.005, 1.85, mask, engine=3) as lat:
        lat.run(steps); lat.sync()
        lat.run(steps); lat.sync()
        ms = lat.last_run_ms()
        pr = lat.profile(0)[:, 0, :].astype(np.float64)       # [task, (poll cycles, extra polls, compute cycles, steps)]
        n = pr[:, 3]
        print(f"{grid}: {ms / steps * 1e3:.2f} us/step = {ms / steps * 1e-3 * 1.965e9:.0f} cycles @1965 MHz | per task and step: "
              f"poll {np.mean(pr[:, 0] / n):.0f} cycles (min {np.min(pr[:, 0] / n):.0f}, max {np.max(pr[:, 0] / n):.0f}), "
              f"extra polls {np.mean(pr[:, 1] / n):.2f}, collide+store {np.mean(pr[:, 2] / n):.0f} cycles "
              f"(min {np.min(pr[:, 2] / n):.0f}, max {np.max(pr[:, 2] / n):.0f})", flush=True)
