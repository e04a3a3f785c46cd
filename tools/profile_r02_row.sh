#!/bin/bash
# ncu --set full of the row-block dataflow kernel (engine 4) on the check grids (run under gpurun on ONE B200).
set -u
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
for grid in 128x128 256x256; do
  E="python tools/explore.py --grid $grid --steps 2000 --reps 1 --configs 4:0:0:1"
  $E > gpurun_out/plain_row_$grid.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:lbm_row -s 1 -c 1 -f -o gpurun_out/prof_row_$grid $E > gpurun_out/ncu_row_$grid.log 2>&1
  echo "row $grid full capture rc=$?"
  tail -1 gpurun_out/plain_row_$grid.log | cut -c1-250
done
ls -la gpurun_out/prof_row_*.ncu-rep
