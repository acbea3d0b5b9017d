#!/bin/sh
# Builds the micro-benchmarks (binaries are git-ignored; they travel to the GPU box with gpurun).
set -e
cd "$(dirname "$0")/.."
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -ccbin /usr/bin/g++"
$NVCC $FLAGS -o tools/ubench tools/ubench.cu
for m in 0 2; do
  $NVCC $FLAGS -DMPCB_FM_MODE=$m -Xptxas -v -o tools/ubench_model_$m tools/ubench_model.cu 2>&1 | grep -E "error|registers|spill" | grep -v " 0 bytes spill" || true
done
