#!/usr/bin/env bash
# round 2, GPU call 1 (one B200): parity tests incl. the large-grid oracle/hash tests and the torn-record stress test,
# engine 3 vs 4 on small lattices, a short default bench (with parity_check, 2048x1024 and CLI legs).
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r02_pytest_gpu.log
echo "== small lattices: engine 3 vs 4"
for g in 128x128 128x256 256x256 512x256 1024x128; do
  timeout 200 python tools/probe_sync.py --synthetic $g --steps 8000 --reps 3 --profile 0 --variants "LBM_ENGINE=3;LBM_ENGINE=4" 2>&1 | grep -v "^ " | cut -c1-200
done | tee gpurun_out/r02_small.log
echo "== bench"; timeout 1200 python bench.py --steps 2 --warmup 3 2> gpurun_out/r02_bench_err.log | tee gpurun_out/r02_bench_n1.json | cut -c1-3000; tail -5 gpurun_out/r02_bench_err.log
