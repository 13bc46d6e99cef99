#!/bin/bash
# Builds a variant of libpiano_step.so into build/<name>.so: tools/build_variant.sh <name> [nvcc flags, e.g. -DPP_WPB=5]
set -e
cd "$(dirname "$0")/../diffusion-piano_b200/csrc"
name=$1; shift
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
COMMON="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
OBJ=$(mktemp -d); trap 'rm -rf "$OBJ"' EXIT
mkdir -p ../../build
$NVCC $COMMON "$@" -c pp_api.cu -o "$OBJ/pp_api.o" &
$NVCC $COMMON "$@" -c pp_piano_only.cu -o "$OBJ/pp_piano_only.o" &
$NVCC $COMMON -prec-div=false -prec-sqrt=false -Xptxas -v "$@" -c pp_hands.cu -o "$OBJ/pp_hands.o" 2> "../../build/$name.ptxas.txt" &
$NVCC $COMMON -prec-div=false -prec-sqrt=false "$@" -c pp_hands_split.cu -o "$OBJ/pp_hands_split.o" &
wait
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o "../../build/$name.so" "$OBJ/pp_api.o" "$OBJ/pp_piano_only.o" "$OBJ/pp_hands.o" "$OBJ/pp_hands_split.o"
grep -A1 "pp_hands_step_kernelILi" "../../build/$name.ptxas.txt" | grep -E "registers|spill" | head -4
