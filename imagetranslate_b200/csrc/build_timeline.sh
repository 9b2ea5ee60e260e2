#!/bin/bash
# Diagnostic build: libitb with per-kernel %globaltimer stamps (see timeline_all.cu).  Output: csrc/build/libitb_timeline.so
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
mkdir -p "$HERE/build"
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -DITB_TIMELINE -shared \
  -o "$HERE/build/libitb_timeline.so" "$HERE/timeline_all.cu"
echo "built $HERE/build/libitb_timeline.so"
