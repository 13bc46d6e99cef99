#!/bin/bash
# Builds libpiano_step.so (sm_100a) in-tree. Usage: csrc/build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -Xcompiler -fPIC -shared -Xptxas -v "$@" \
  -o ../libpiano_step.so pp_api.cu pp_piano_only.cu pp_hands.cu 2>&1
