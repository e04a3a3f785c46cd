#!/usr/bin/env bash
# lean 8-GPU confirmation: the multi-GPU parity tests (all engines, both process models, checked build) and the default
# torchrun bench at N = 8
set -u
mkdir -p gpurun_out
echo "== pytest multi (8 GPUs)"; timeout 900 python -m pytest tests -m gpu -x -q -k "slabs or one_process or torn or checked" --durations=4 2>&1 | tail -10 | tee gpurun_out/r02_pytest_multi_8.log
echo "== bench N=8"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus 8 --steps 3 --warmup 3 2> gpurun_out/r02_bench_err_8.log | tee gpurun_out/r02_bench_n8_final.json | cut -c1-300
grep -i "error\|Traceback" gpurun_out/r02_bench_err_8.log | head -5
