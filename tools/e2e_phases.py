#!/usr/bin/env python
"""Prints bench.py's e2e phase times (create / run / download / destroy) for one grid, step by step.
usage: python tools/e2e_phases.py 1024x1024 [iters] [steps]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

lbm = importlib.import_module("lattice-boltzmann_b200")
grid = sys.argv[1] if len(sys.argv) > 1 else "1024x1024"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
torch.cuda.set_device(0)
for i in range(steps):
    r = bench.time_e2e(lbm, torch, grid, iters, 1, 0, 1, None)
    print(grid, iters, i, round(r["glups"], 1), r["phases"], flush=True)
