#!/usr/bin/env bash
# One GPU-box round: tests, smoke, bench, and the ncu evidence for profiles/.
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee gpurun_out/smoke.log
echo "== bench"; timeout 900 python bench.py --steps 3 --warmup 3 2> gpurun_out/bench_err.log | tee gpurun_out/bench_n1.json; tail -3 gpurun_out/bench_err.log
echo "== bench reference arm"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 --ref-seconds 8 2>&1 | tee gpurun_out/bench_ref.json
CMD="python tools/explore.py --grid 1024x1024 --steps 300 --reps 1 --configs 1:256:2:1"
echo "== ncu launch list"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/plain.log
echo "== ncu full"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:lbm_persistent -c 1 -o gpurun_out/prof_persistent_1024 -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/ncu_full.log
