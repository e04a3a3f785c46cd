#!/usr/bin/env bash
# Multi-GPU round: N-GPU == 1-GPU parity, then torchrun bench at N GPUs.
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m 2>&1 | head -12
echo "== pytest multi"; timeout 900 python -m pytest tests -m gpu -q -k "slabs or two_gpu" 2>&1 | tail -8 | tee gpurun_out/pytest_multi_$N.log
echo "== explore single-process N=$N"; timeout 300 python tools/explore.py --grid 16384x1024 --steps 300 --reps 2 --ngpus $N --configs 1:256:2:1
echo "== torchrun bench N=$N"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 2 --warmup 3 --iters 2000 2> gpurun_out/bench_multi_err_$N.log | tee gpurun_out/bench_n$N.json
tail -5 gpurun_out/bench_multi_err_$N.log
