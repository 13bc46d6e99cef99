#!/bin/bash
# Builds libpiano_step.so and libppo_kernels.so (sm_100a) in-tree. Usage: csrc/build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
# pp_hands.cu only: -prec-div=false -prec-sqrt=false (2-ulp division / square root: +4.7 % env-steps/s, parity
# against the fp64 restatement unchanged). The piano-only kernel keeps IEEE division (its closed-form test is tight),
# and nothing uses --use_fast_math (approximate logf breaks the cross-entropy reward test).
COMMON="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v"
OBJ=$(mktemp -d)
trap 'rm -rf "$OBJ"' EXIT
$NVCC $COMMON "$@" -c pp_api.cu -o "$OBJ/pp_api.o" 2>&1
$NVCC $COMMON "$@" -c pp_piano_only.cu -o "$OBJ/pp_piano_only.o" 2>&1
$NVCC $COMMON -prec-div=false -prec-sqrt=false "$@" -c pp_hands.cu -o "$OBJ/pp_hands.o" 2>&1
$NVCC $COMMON -prec-div=false -prec-sqrt=false "$@" -c pp_hands_split.cu -o "$OBJ/pp_hands_split.o" 2>&1
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../libpiano_step.so "$OBJ/pp_api.o" "$OBJ/pp_piano_only.o" "$OBJ/pp_hands.o" "$OBJ/pp_hands_split.o" 2>&1
$NVCC $COMMON "$@" -shared ppo_kernels.cu -o ../libppo_kernels.so 2>&1
