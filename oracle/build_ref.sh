#!/usr/bin/env bash
# Builds the UNMODIFIED reference (d2q9-bgk.c) into oracle/_ref/ as the CPU oracle
# and timed CPU baseline. TEST INFRASTRUCTURE: nothing under oracle/ is linked into
# or called by the product (liblbm_b200.so / d2q9-bgk).
#
#   ref-serial : gcc -O3, exactly the file as shipped (pragmas commented out)
#   ref-omp    : the reference's OpenMP variant = the same file with its three
#                commented-out "#pragma omp parallel for" lines (d2q9-bgk.c:278,588,607)
#                re-enabled on the fly by sed, + -fopenmp (reference Makefile:7)
#   ref-debug  : ref-serial with -DDEBUG (per-step "av velocity" / "tot density" lines, :198-202)
#
# The source is read where it lies under $REF (default /root/reference); no copy of
# it is written anywhere (the sed output is piped straight into gcc).  Outputs go
# only into oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun).
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REF:-/root/reference}"
out="$here/_ref"
src="$REF/d2q9-bgk.c"
if [ ! -f "$src" ]; then
    echo "build_ref.sh: $src not found; keeping prebuilt oracle/_ref (if any)" >&2
    exit 0
fi
mkdir -p "$out"
# Intel-only intrinsics in the reference (d2q9-bgk.c:291-312) become no-ops under gcc.
DEFS=(-D'__assume(x)=((void)0)' -D'__assume_aligned(p,a)=((void)0)')
CFLAGS=(-std=gnu99 -O3 -I"$here/mpi_stub" -w)
gcc "${CFLAGS[@]}" "${DEFS[@]}" "$src" -lm -o "$out/ref-serial"
gcc "${CFLAGS[@]}" "${DEFS[@]}" -DDEBUG "$src" -lm -o "$out/ref-debug"
sed 's#^\( *\)// \#pragma omp parallel for#\1\#pragma omp parallel for#' "$src" \
  | gcc "${CFLAGS[@]}" "${DEFS[@]}" -fopenmp -x c - -lm -o "$out/ref-omp"
echo "built $out/ref-serial $out/ref-omp $out/ref-debug"
