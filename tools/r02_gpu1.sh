#!/usr/bin/env bash
# round 2 (one B200): parity tests incl. the large-grid oracle/hash tests and the torn-record stress test, then a short
# default bench (with parity_check, small-lattice, 2048x1024 and CLI legs).
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 2>&1 | tail -25 | tee gpurun_out/r02_pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
echo "== bench"; timeout 1200 python bench.py --steps 2 --warmup 3 2> gpurun_out/r02_bench_err.log | tee gpurun_out/r02_bench_n1.json | cut -c1-6000; tail -5 gpurun_out/r02_bench_err.log
