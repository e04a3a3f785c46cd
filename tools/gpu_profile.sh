#!/usr/bin/env bash
# ncu evidence for profiles/: launch list of the bench command + full captures of the timestep kernel.
set -u
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --iters 100 --no-cpu-baseline"
echo "== launch list of: $B"
$B > gpurun_out/plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench.csv $B > gpurun_out/ncu_list.log 2>&1
echo "rc=$?"
for G in 1024x1024 16384x1024; do
  S=300; [ $G = 16384x1024 ] && S=40
  CMD="python tools/explore.py --grid $G --steps $S --reps 1 --configs 1:256:2:1"
  echo "== ncu full $G"
  $CMD > gpurun_out/plain_$G.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:lbm_persistent -s 1 -c 1 -f -o gpurun_out/prof_persistent_$G $CMD > gpurun_out/ncu_full_$G.log 2>&1
  echo "rc=$?"; tail -2 gpurun_out/plain_$G.log
done
