#!/bin/bash
# Builds imagetranslate_b200/libitb.so for sm_100a (cross-compiles without a GPU).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../libitb.so"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC ${ITB_NVCC_EXTRA}"
mkdir -p "$HERE/build"
pids=()
for f in misc attention attention_mma attention_bulk beam gemm_simt gemm_tc api no_alloc; do
  if [ ! -f "$HERE/build/$f.o" ] || [ "$HERE/$f.cu" -nt "$HERE/build/$f.o" ] || [ "$HERE/common.cuh" -nt "$HERE/build/$f.o" ] || [ "$HERE/launch.h" -nt "$HERE/build/$f.o" ] || [ "$HERE/mma16.cuh" -nt "$HERE/build/$f.o" ] || [ "$HERE/../../include/itb.h" -nt "$HERE/build/$f.o" ]; then
    $NVCC $FLAGS -c "$HERE/$f.cu" -o "$HERE/build/$f.o" &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o "$OUT" "$HERE"/build/{misc,attention,attention_mma,attention_bulk,beam,gemm_simt,gemm_tc,api,no_alloc}.o
echo "built $OUT"
