#!/usr/bin/env bash
# small lattices: engines 3 and 4 side by side (device-timed, 8000 timesteps per launch, best of 3)
set -u
mkdir -p gpurun_out
for g in 128x128 128x256 256x256 512x256 1024x128 512x512; do
  timeout 300 python tools/probe_sync.py --synthetic $g --steps 8000 --reps 3 --profile 0 --variants "LBM_ENGINE=3;LBM_ENGINE=4" 2>&1 | grep -v "^ " | sed 's/ctas\/slab.*interior=[0-9]*%\]//' | cut -c1-200
done | tee gpurun_out/r02_small.log
