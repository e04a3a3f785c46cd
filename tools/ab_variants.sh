#!/usr/bin/env bash
# A/B of separately built libraries inside ONE gpurun call (same box, same clocks) - how the kernel variants in
# profiles/r01_small_lattice.md and r01_sync_probe.md were compared.
#   here:      make && cp lattice-boltzmann_b200/liblbm_b200.so scratch/a.so   (edit, rebuild) ... scratch/b.so
#   GPU box:   gpurun -- 'bash tools/ab_variants.sh "--synthetic 256x256 --steps 8000 --variants LBM_ENGINE=3" scratch/a.so scratch/b.so'
# Every variant is run twice, interleaved (a b a b), so drift shows up as disagreement between the two passes.
# scratch/ is git-ignored but travels to the GPU box.
set -u
ARGS=$1; shift
LIB=lattice-boltzmann_b200/liblbm_b200.so
cp $LIB /tmp/ab_keep.so
for pass in 1 2; do
  for v in "$@"; do
    cp "$v" $LIB
    echo -n "$(basename "$v" .so) pass $pass: "
    timeout 300 python tools/probe_sync.py --ngpus 1 --profile 0 --reps 3 $ARGS 2>&1 | tail -1 | grep -o "[0-9.]* us/step [0-9.]* GLUPS.*"
  done
done
cp /tmp/ab_keep.so $LIB
