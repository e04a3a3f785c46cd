#!/bin/bash
# ncu evidence for profiles/ (run under gpurun on ONE B200):
#   1. launch list of a short default-workload bench run (per-launch device time; compare SHARES)
#   2. ncu --set full of the persistent timestep kernel on 16384x1024 (HBM-bound) and 1024x1024 (L2-resident),
#      and of the small-lattice dataflow kernel on 256x256
# Every command runs plain first and is profiled only if that exits 0.
set -u
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --iters 300 --no-cpu-baseline --no-also"
$B > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench.csv $B > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
for G in 16384x1024:40 1024x1024:300; do
  grid=${G%%:*}; steps=${G##*:}
  E="python tools/explore.py --grid $grid --steps $steps --reps 1 --configs 1:256:2:1"
  $E > gpurun_out/plain_$grid.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:lbm_persistent -s 1 -c 1 -f -o gpurun_out/prof_persistent_$grid $E > gpurun_out/ncu_$grid.log 2>&1
  echo "$grid full capture rc=$?"
  tail -2 gpurun_out/plain_$grid.log
done
# engine 3 (dataflow over tagged records) on 256x256
E="python tools/explore.py --grid 256x256 --steps 2000 --reps 1 --configs 3:32:0:1"
$E > gpurun_out/plain_df_256x256.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:lbm_dataflow -s 1 -c 1 -f -o gpurun_out/prof_dataflow_256x256 $E > gpurun_out/ncu_df_256x256.log 2>&1
echo "dataflow 256x256 full capture rc=$?"
tail -2 gpurun_out/plain_df_256x256.log
ls -la gpurun_out/*.ncu-rep
