#!/usr/bin/env bash
# compute-sanitizer over the drop-in program on the small check grids (SURVEY.md section 5 "race detection": the
# reference has none).  Every engine, every tool; a 2-GPU row-slab case when two devices are visible.  The kernels
# synchronise through hand-rolled release/acquire flags, spin loops and tagged records, so memcheck (out-of-bounds,
# misaligned), initcheck (reads of never-written device memory), synccheck (barrier misuse) and racecheck (shared
# memory hazards) each have something to say.  Runs on the GPU box:  gpurun -- 'bash tools/sanitize.sh'
# Output: gpurun_out/sanitize.log (full) and a one-line verdict per case on stdout.
set -u
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/gpurun_out; mkdir -p "$OUT"; LOG=$OUT/sanitize.log; : > "$LOG"
W=$(mktemp -d); cd "$W"
STEPS=${SANITIZE_STEPS:-24}
NG=$(nvidia-smi -L 2>/dev/null | wc -l)
fail=0
run_case() {   # name, tool, env..., -- grid
  local name=$1 tool=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  local grid=$1
  sed "3s/.*/$STEPS/" "$ROOT/inputs/input_$grid.params" > p.params
  echo "=== $name: $tool ${envs[*]} $grid ($STEPS steps)" >> "$LOG"
  env "${envs[@]}" timeout 600 compute-sanitizer --tool "$tool" --error-exitcode 86 --print-limit 20 \
      "$ROOT/d2q9-bgk" p.params "$ROOT/inputs/obstacles_$grid.dat" >> "$LOG" 2>&1
  local rc=$?
  local errs; errs=$(grep -c "^========= .*\(Invalid\|Uninitialized\|hazard\|Barrier error\|error\)" "$LOG" || true)
  if [ $rc -eq 0 ]; then echo "PASS $name $tool $grid"; else echo "FAIL($rc) $name $tool $grid"; fail=1; fi
}
for tool in memcheck initcheck synccheck racecheck; do
  for eng in 1 2 3 4; do
    for grid in 128x128 128x256; do
      run_case "engine$eng" $tool LBM_ENGINE=$eng -- $grid
    done
  done
  if [ "$NG" -ge 2 ]; then
    run_case "2gpu-ring" $tool LBM_NGPUS=2 -- 128x256
    run_case "2gpu-ring-dataflow" $tool LBM_NGPUS=2 LBM_ENGINE=3 -- 128x256
  fi
done
grep -E "ERROR SUMMARY|^===" "$LOG" | paste - - | sed 's/=========//' > "$OUT/sanitize_summary.txt"
cat "$OUT/sanitize_summary.txt" | tail -60
exit $fail
