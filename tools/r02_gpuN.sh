#!/usr/bin/env bash
# round 2, N GPUs of one box: the multi-GPU parity tests (engine 1 halo slots and the dataflow engines across GPUs, both
# process models), the torn-record stress test over NVLink, then a short torchrun bench with parity_check.
set -u
N=${1:-2}
ITERS=${2:-2000}
mkdir -p gpurun_out
echo "== pytest multi"; timeout 1500 python -m pytest tests -m gpu -x -q -k "torn or slabs or one_process" --durations=5 2>&1 | tail -22 | tee gpurun_out/r02_pytest_multi_$N.log
echo "== torchrun bench N=$N"
NCCL_DEBUG=VERSION timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 2 --warmup 3 --iters $ITERS 2> gpurun_out/r02_bench_err_$N.log | tee gpurun_out/r02_bench_n$N.json | cut -c1-5000
grep -i "error\|Traceback\|NCCL version" gpurun_out/r02_bench_err_$N.log | head -5
