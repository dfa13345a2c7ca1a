#!/bin/bash
# Builds the two micro-benchmarks (tools/tma_probe, tools/mma_probe) for sm_100a; no GPU needed to build.
set -e
cd "$(dirname "$0")/.."
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
for p in tma_probe mma_probe; do
  $NVCC -gencode arch=compute_100a,code=sm_100a -O3 -I. tools/$p.cu -o tools/$p -lcuda
  echo "built tools/$p"
done
