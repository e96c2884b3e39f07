#!/bin/bash
# Builds libb2llm.so in-tree for sm_100a (cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
OUT=${OUT:-../libb2llm.so}
BUILD=${BUILD:-build}
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fopenmp,-ffp-contract=off,-fvisibility=default -cudart static"
mkdir -p $BUILD
pids=()
for f in kernels.cu gemv.cu gemv_tp.cu attn.cu prefill.cu engine.cu ops.cu sampler.cu noalloc.cu; do
  $NVCC $FLAGS ${EXTRA:-} -c $f -o $BUILD/${f%.cu}.o &
  pids+=($!)
done
g++ -O2 -std=c++17 -fPIC -c nvml.cpp -o $BUILD/nvml.o &
pids+=($!)
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared $FLAGS -o $OUT $BUILD/kernels.o $BUILD/gemv.o $BUILD/gemv_tp.o $BUILD/attn.o $BUILD/prefill.o $BUILD/engine.o $BUILD/ops.o $BUILD/sampler.o $BUILD/noalloc.o $BUILD/nvml.o -lgomp -ldl
echo "built $(realpath $OUT)"
