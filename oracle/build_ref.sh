#!/bin/bash
# TEST INFRASTRUCTURE. Builds the UNMODIFIED reference sources (where they lie under
# /root/reference/MTPC) + oracle/ref_driver.cpp into oracle/_ref/mcpt_ref, and makes LF copies of
# the scene files into oracle/_ref/scene/ (the reference parsers keep '\r' on Linux).
# Outputs go only to oracle/_ref/ (git-ignored, travels to the GPU box). No reference source is copied.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${MCPT_REFERENCE:-/root/reference}/MTPC"
OUT="$HERE/_ref"
mkdir -p "$OUT/obj" "$OUT/scene" "$OUT/result" "$OUT/cwd"
if [ ! -d "$REF" ]; then echo "reference not present at $REF; keeping prebuilt $OUT" >&2; exit 0; fi
CXX="${MCPT_CXX:-/usr/bin/g++}"   # not $CXX: this image exports a wrapper that cannot find libgomp.spec
FLAGS="-O3 -std=c++17 -fopenmp -w -I$HERE/shim -I$REF -include $HERE/shim/winshim.h"
for src in "morton code" BVH sceneManagement pathTracing; do
  $CXX $FLAGS -c "$REF/$src.cpp" -o "$OUT/obj/${src// /_}.o"
done
$CXX $FLAGS -Dmain=ref_main -c "$REF/MTPC.cpp" -o "$OUT/obj/MTPC.o"
$CXX $FLAGS -c "$HERE/ref_driver.cpp" -o "$OUT/obj/ref_driver.o"
$CXX -O3 -fopenmp "$OUT"/obj/*.o -o "$OUT/mcpt_ref"
for f in "$HERE"/../scenes/*.obj "$HERE"/../scenes/*.mtl "$HERE"/../scenes/*.camera; do
  tr -d '\r' < "$f" > "$OUT/scene/$(basename "$f")"
done
cp "$HERE"/../scenes/*.bgr "$OUT/cwd/"
echo "built $OUT/mcpt_ref"
