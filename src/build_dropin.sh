#!/usr/bin/env bash
# src/build_dropin.sh — builds the kzg.h drop-in: the reference's UNMODIFIED headers and its unmodified
# blob/commit/proof/util sources (compiled where they lie under /root/reference) + OUR
# src/trusted_setup_gpu.cpp and src/poly_gpu.cpp in place of the reference's src/trusted_setup.cpp and src/poly.cpp, linked against
# kzg-commitments_b200/libkzg_gpu.so.  Then links the reference's own callers against it, unchanged:
#   oracle/_ref/dropin/testing-bn254-gpu        testing/testing.cpp   (36 property checks)
#   oracle/_ref/dropin/benchmark-<curve>-gpu    benchmark/benchmark.cpp
#   oracle/_ref/dropin/kzg-cli-bn254-gpu        demo/shared/kzg-cli.cpp (the demo's setup / commit / prove / verify CLI)
#   oracle/_ref/dropin/kzgdrop-<curve>          oracle/ref_harness.cpp (same byte-file driver as kzgref-<curve>)
#   oracle/_ref/dropin/dropin-threads-bn254     tests/cpp/dropin_threads.cpp (concurrent callers, cache eviction, copies)
# Needs oracle/build_ref.sh to have run (MIRACL core.a, NTL ntl.a and generated headers in oracle/_ref/scratch).
# Outputs are git-ignored and travel to the GPU box like the other built artefacts; /root/reference is
# only needed here, at build time.
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(cd "$HERE/.." && pwd)
SCR=$ROOT/oracle/_ref/scratch
OUT=$ROOT/oracle/_ref/dropin
PKG=$ROOT/kzg-commitments_b200
if [ ! -d "$REF/src" ] || [ ! -f "$SCR/miracl/core.a" ]; then
  echo "build_dropin: reference or oracle/_ref/scratch not present; keeping prebuilt $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"
build_curve() {
  local c=$1 inc=$SCR/include-$1 obj=$SCR/dropin-obj-$1
  mkdir -p "$obj"
  local I="-I$inc -I$SCR/ntl/include -I$REF/src"
  for f in blob commit proof util; do
    g++ -O2 $I -c "$REF/src/$f.cpp" -o "$obj/$f.o"
  done
  rm -f "$obj/poly.o"
  g++ -O2 -std=c++17 $I -c "$HERE/trusted_setup_gpu.cpp" -o "$obj/trusted_setup_gpu.o"
  g++ -O2 -std=c++17 $I -c "$HERE/poly_gpu.cpp" -o "$obj/poly_gpu.o"
  rm -f "$OUT/libkzg-$c-gpu.a"; ar rcs "$OUT/libkzg-$c-gpu.a" "$obj"/*.o
  local L="$OUT/libkzg-$c-gpu.a $SCR/miracl/core.a $SCR/ntl/src/ntl.a -L$PKG -lkzg_gpu -lpthread -Wl,-rpath,\$ORIGIN/../../../kzg-commitments_b200"
  g++ -O2 $I "$ROOT/oracle/ref_harness.cpp" $L -o "$OUT/kzgdrop-$c"
  g++ -O2 $I "$REF/benchmark/benchmark.cpp" $L -o "$OUT/benchmark-$c-gpu"
  if [ "$c" = bn254 ]; then
    g++ -O2 $I "$REF/testing/testing.cpp" $L -o "$OUT/testing-bn254-gpu"
    g++ -O2 $I "$REF/demo/shared/kzg-cli.cpp" $L -o "$OUT/kzg-cli-bn254-gpu"
    g++ -O2 -std=c++17 $I "$ROOT/tests/cpp/dropin_threads.cpp" $L -o "$OUT/dropin-threads-bn254"   # our own thread-safety test
    cp "$REF/testing/blob1.txt" "$REF/testing/blob2.txt" "$OUT/" 2>/dev/null || true
  fi
}
build_curve bls12381 &
build_curve bn254 &
build_curve bn158 &
wait
ls -la "$OUT"
