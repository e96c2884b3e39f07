#!/bin/bash
# Builds libb2llm.so in-tree for sm_100a (cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
OUT=../libb2llm.so
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fopenmp,-ffp-contract=off,-fvisibility=default -cudart static"
mkdir -p build
pids=()
for f in kernels.cu gemv.cu gemv_tp.cu attn.cu mega.cu prefill.cu engine.cu ops.cu sampler.cu noalloc.cu; do
  $NVCC $FLAGS ${EXTRA:-} -c $f -o build/${f%.cu}.o &
  pids+=($!)
done
g++ -O2 -std=c++17 -fPIC -c nvml.cpp -o build/nvml.o &
pids+=($!)
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared $FLAGS -o $OUT build/kernels.o build/gemv.o build/gemv_tp.o build/attn.o build/mega.o build/prefill.o build/engine.o build/ops.o build/sampler.o build/noalloc.o build/nvml.o -lgomp -ldl
echo "built $(realpath $OUT)"
