// gemv_launch.cuh — the stand-alone GEMV kernel and its launch templates, shared by gemv.cu (single-GPU modes) and
// gemv_tp.cu (the tensor-parallel input / output modes, compiled as a translation unit of their own).
#pragma once
#include "gemv_dev.cuh"

// ------------------------------------------------------------------------------------------ the kernel
template <int IN, int OUT, int PATH, int QT, class A>
__global__ void __launch_bounds__(GEMV_THREADS, 1) k_gemv(const A a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint32_t parity = 0;
    const GemvPtrs pv = {a.w, a.norm_w, a.bias_or_res};
    gemv_phase<IN, OUT, PATH, QT, A>(a, pv, smem, parity);
}

template <int IN, int OUT, int PATH, int QT, class A>
static cudaError_t launch_path(const A &a, int grid, size_t smem, cudaStream_t st, bool pdl) {
    // function attributes belong to a device: a server process drives one engine per GPU (server.py --devices)
    static bool attr_set[B2_MAX_DEVICES] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= B2_MAX_DEVICES || !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(k_gemv<IN, OUT, PATH, QT, A>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < B2_MAX_DEVICES) attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GEMV_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, k_gemv<IN, OUT, PATH, QT, A>, a);
}
template <int IN, int OUT, int QT, class A>
static cudaError_t launch_qt(const A &a, int grid, size_t smem, cudaStream_t st, bool pdl) {
    const int cpt = (a.nbp + KC - 1) / KC;
    if (a.k_outer) return launch_path<IN, OUT, 2, QT>(a, grid, smem, st, pdl);
    if (cpt == 2) return launch_path<IN, OUT, 0, QT>(a, grid, smem, st, pdl);
    return launch_path<IN, OUT, 1, QT>(a, grid, smem, st, pdl);
}
template <int IN, int OUT, class A>
static cudaError_t launch_one(const A &a, int grid, size_t smem, cudaStream_t st, bool pdl) {
    if (a.qt == 2) return launch_qt<IN, OUT, 2>(a, grid, smem, st, pdl);
    return a.qt ? launch_qt<IN, OUT, 1>(a, grid, smem, st, pdl) : launch_qt<IN, OUT, 0>(a, grid, smem, st, pdl);
}

