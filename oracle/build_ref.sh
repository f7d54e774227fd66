#!/usr/bin/env bash
# oracle/build_ref.sh — builds the UNMODIFIED reference (MIRACL Core + NTL + src/) from
# /root/reference into oracle/_ref/ (git-ignored, travels to the GPU box).  TEST INFRASTRUCTURE.
#
# Outputs (per curve c in bls12381 bn254 bn158):
#   oracle/_ref/kzgref-$c        our harness (oracle/ref_harness.cpp) linked against the reference
#   oracle/_ref/benchmark-$c     the reference's own benchmark/benchmark.cpp, unmodified
#   oracle/_ref/testing-bn254    the reference's own testing/testing.cpp, unmodified
#   oracle/_ref/kzg-cli-bn254    the reference's demo CLI demo/shared/kzg-cli.cpp, unmodified
# Nothing from /root/reference is copied into tracked paths; the scratch copy lives in
# oracle/_ref/scratch (listed in .gpurunignore).
#
# Recipe = SURVEY.md §8(c): MIRACL `config64.py -o 28 -o 31 -o 41` (its template expansion is the
# only way to obtain the per-curve sources), NTL `./configure NTL_GMP_LIP=off` (gmp.h is absent here).
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
SCR=$OUT/scratch
JOBS=${JOBS:-8}

if [ ! -d "$REF/src" ]; then
  echo "build_ref: $REF not present; keeping prebuilt $OUT" >&2
  exit 0
fi
mkdir -p "$SCR"

if [ ! -f "$SCR/miracl/core.a" ]; then
  rm -rf "$SCR/miracl"; cp -r "$REF/miracl-core/cpp" "$SCR/miracl"; chmod -R u+w "$SCR/miracl"
  (cd "$SCR/miracl" && python3 config64.py -o 28 -o 31 -o 41 > "$SCR/miracl.log" 2>&1)
fi
if [ ! -f "$SCR/ntl/src/ntl.a" ]; then
  rm -rf "$SCR/ntl"; cp -r "$REF/ntl" "$SCR/ntl"; chmod -R u+w "$SCR/ntl"
  (cd "$SCR/ntl/src" && ./configure NTL_GMP_LIP=off > "$SCR/ntl_configure.log" 2>&1 && make -j"$JOBS" > "$SCR/ntl_make.log" 2>&1)
fi

build_curve() {  # $1=lower name  $2=MIRACL curve tag  $3=BIG tag
  local c=$1 tag=$2 big=$3 inc=$SCR/include-$1 obj=$SCR/obj-$1
  mkdir -p "$inc" "$obj"
  cp "$SCR"/miracl/*_"$tag".h "$SCR"/miracl/*_"$big".h "$SCR"/miracl/arch.h "$SCR"/miracl/core.h "$SCR"/miracl/randapi.h "$inc"/
  cp "$REF/config/curve_$tag/kzg_config.h" "$inc/kzg_config.h"
  cp "$REF/src/kzg.h" "$inc/kzg.h"
  local I="-I$inc -I$SCR/ntl/include -I$REF/src"
  for f in "$REF"/src/*.cpp; do
    g++ -O2 $I -c "$f" -o "$obj/$(basename "$f" .cpp).o"
  done
  rm -f "$SCR/kzg-$c.a"; ar rcs "$SCR/kzg-$c.a" "$obj"/*.o
  local L="$SCR/kzg-$c.a $SCR/miracl/core.a $SCR/ntl/src/ntl.a -lpthread"
  g++ -O2 $I "$HERE/ref_harness.cpp" $L -o "$OUT/kzgref-$c"
  g++ -O2 $I "$REF/benchmark/benchmark.cpp" $L -o "$OUT/benchmark-$c"
  if [ "$c" = bn254 ]; then
    g++ -O2 $I "$REF/testing/testing.cpp" $L -o "$OUT/testing-bn254"
    g++ -O2 $I "$REF/demo/shared/kzg-cli.cpp" $L -o "$OUT/kzg-cli-bn254"   # the demo CLI (reference Makefile:14-15)
  fi
}
build_curve bls12381 BLS12381 B384_58 &
build_curve bn254 BN254 B256_56 &
build_curve bn158 BN158 B160_56 &
wait
ls -la "$OUT" | grep -v scratch
