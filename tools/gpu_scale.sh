#!/usr/bin/env bash
# Strong-scaling series on one box: parity at every slab count, then torchrun bench at N = 2,4,8 (and N=1).
set -u
MAXN=${1:-8}
mkdir -p gpurun_out
echo "== pytest multi"; timeout 900 python -m pytest tests -m gpu -q -k "slabs or two_gpu" 2>&1 | tail -6 | tee gpurun_out/pytest_multi_$MAXN.log
for N in 1 2 4 8; do
  [ $N -gt $MAXN ] && break
  echo "== single-process explore N=$N"; timeout 300 python tools/explore.py --grid 16384x1024 --steps 400 --reps 2 --ngpus $N --configs 1:256:2:1 2>&1 | cut -c1-40,170-300
done
for N in 2 4 8; do
  [ $N -gt $MAXN ] && break
  echo "== torchrun bench N=$N"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 2 --warmup 3 --iters 4000 2> gpurun_out/bench_multi_err_$N.log | tee gpurun_out/bench_n$N.json | cut -c1-220
  grep -i "error\|Traceback" gpurun_out/bench_multi_err_$N.log | head -5
done
echo "== CLI LBM_NGPUS=$MAXN on 4096x1024 (reduced iters)"; mkdir -p /tmp/cli && cd /tmp/cli && sed '3s/.*/2000/' $OLDPWD/inputs/input_4096x1024.params > p.params && LBM_REPORT=1 LBM_NGPUS=$MAXN timeout 300 $OLDPWD/d2q9-bgk p.params $OLDPWD/inputs/obstacles_4096x1024.dat 2>&1 | tail -12
