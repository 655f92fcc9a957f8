#!/bin/bash
# Builds the library of another git revision next to the working tree's, for A/B runs on one box:
#   tools/ab_build.sh <rev> [name]   ->  build/ab/<name>/libcgb200.so
#   CGB200_LIBRARY=build/ab/<name>/libcgb200.so python tools/dense_trace.py
set -e
REV=${1:?usage: ab_build.sh <rev> [name]}
NAME=${2:-$REV}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/build/ab/$NAME
TMP=$(mktemp -d)
git -C "$ROOT" archive "$REV" conjugate-gradient_b200 include | tar -x -C "$TMP"
make -C "$TMP/conjugate-gradient_b200" libcgb200.so -j4 > "$TMP/build.log" 2>&1 || { tail -20 "$TMP/build.log"; exit 1; }
mkdir -p "$OUT"
cp "$TMP/conjugate-gradient_b200/libcgb200.so" "$OUT/"
rm -rf "$TMP"
echo "$OUT/libcgb200.so"
