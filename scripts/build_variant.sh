#!/bin/bash
# Build a tuning variant of the SAME sources: scripts/build_variant.sh NAME "<extra nvcc flags>"
#   -> tfhe_b200/variants/libluxtfhe_b200_NAME.so   (select with LUX_TFHE_LIB=<that path>)
# Used only for A/B measurements on the GPU box; the shipped library is tfhe_b200/libluxtfhe_b200.so.
set -eu
NAME=$1; shift
EXTRA="${1:-}"
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=$ROOT/tfhe_b200/csrc
OBJ=$(mktemp -d)
OUT=$ROOT/tfhe_b200/variants
mkdir -p "$OUT"
ARCH="-gencode arch=compute_100a,code=sm_100a"
pids=()
for f in engine radix br_n9 br_n10 br_n11 br_n12; do
  nvcc -O3 -std=c++17 $ARCH -lineinfo -Xcompiler -fPIC $EXTRA -c -o "$OBJ/$f.o" "$SRC/$f.cu" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
nvcc $ARCH -shared -o "$OUT/libluxtfhe_b200_$NAME.so" "$OBJ"/*.o
rm -rf "$OBJ"
echo "$OUT/libluxtfhe_b200_$NAME.so"
