#!/usr/bin/env bash
# round 2, all 8 GPUs of one box: every -m gpu test (multi-GPU ones included), the default torchrun bench at N = 8
# (parity_check + the few-rows-per-GPU legs), the same small lattice forced onto engine 1 for comparison, and N = 4, 2, 1.
set -u
mkdir -p gpurun_out
if [ "${SKIP_TESTS:-0}" != 1 ]; then
echo "== pytest -m gpu (8 GPUs)"; timeout 1500 python -m pytest tests -m gpu -x -q --durations=6 2>&1 | tail -14 | tee gpurun_out/r02_pytest_gpu8.log
fi
run_bench() { # N, extra args...
  local N=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N "$@"
}
echo "== bench N=8 (default)"; NCCL_DEBUG=INFO run_bench 8 --steps 3 --warmup 3 2> gpurun_out/r02_bench_err_8.log | tee gpurun_out/r02_bench_n8.json | cut -c1-300
grep -i "error\|Traceback" gpurun_out/r02_bench_err_8.log | head -5; grep -c "NCCL INFO" gpurun_out/r02_bench_err_8.log
echo "== 1024x1024 on 8 GPUs, engine 1 forced"; LBM_ENGINE=1 run_bench 8 --steps 3 --warmup 3 --workload 1024x1024 --no-also --no-parity 2>> gpurun_out/r02_bench_err_8.log | tee gpurun_out/r02_bench_n8_1024_engine1.json | cut -c1-200
echo "== 131072x128 on 8 GPUs, dataflow engine forced"; LBM_ENGINE=3 run_bench 8 --steps 2 --warmup 3 --workload 131072x128 --iters 1000 --no-also --no-parity 2>> gpurun_out/r02_bench_err_8.log | tee gpurun_out/r02_bench_n8_131072_engine3.json | cut -c1-200
for N in 4 2; do
  echo "== bench N=$N"; run_bench $N --steps 2 --warmup 3 2> gpurun_out/r02_bench_err_$N.log | tee gpurun_out/r02_bench_n$N.json | cut -c1-200
done
echo "== bench N=1"; timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2> gpurun_out/r02_bench_err_1.log | tee gpurun_out/r02_bench_n1_8gpubox.json | cut -c1-200
