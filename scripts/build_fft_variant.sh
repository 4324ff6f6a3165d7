#!/bin/bash
# Tuning variant of the torus kernels only: scripts/build_fft_variant.sh NAME "<extra nvcc flags>"
#   recompiles br_fft.cu with the flags and links it with the objects of the last `make` ->
#   tfhe_b200/variants/libluxtfhe_b200_NAME.so   (select with LUX_TFHE_LIB=<that path>)
set -eu
NAME=$1; shift
EXTRA="${1:-}"
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=$ROOT/tfhe_b200/csrc
OUT=$ROOT/tfhe_b200/variants
mkdir -p "$OUT"
ARCH="-gencode arch=compute_100a,code=sm_100a"
nvcc -O3 -std=c++17 $ARCH -lineinfo -Xcompiler -fPIC $EXTRA -c -o "$OUT/br_fft_$NAME.o" "$SRC/br_fft.cu"
nvcc $ARCH -shared -o "$OUT/libluxtfhe_b200_$NAME.so" "$OUT/br_fft_$NAME.o" "$SRC"/{engine,radix,multi,serial,br_n9,br_n10,br_n11,br_n12}.o
echo "$OUT/libluxtfhe_b200_$NAME.so"
