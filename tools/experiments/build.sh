#!/bin/bash
# Builds tools/experiments/libppo_experiments.so: the product PPO kernels plus the measured-slower experiments
# (ppo_fusion_experiments.cu). Not part of build(); only the probes under tools/ load it.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared \
  ppo_fusion_experiments.cu -o libppo_experiments.so "$@"
