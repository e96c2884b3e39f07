// microbench4.cu — does the ASSIGNMENT of a weight matrix to the SMs matter for HBM read bandwidth?
// 148 CTAs x 512 threads stream 302 MB with 128-bit loads, 4 loads in flight per thread, as the GEMV does:
//   A  every CTA owns one contiguous 2 MB range (the GEMV's layout today)
//   B  36 KB tiles dealt round-robin: CTA b reads tiles b, b+148, ... (all SMs inside a moving ~5 MB window)
//   C  as B with 18 KB chunks
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/microbench4.cu -o tools/bin/microbench4
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint4 ldg4(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// unit = bytes dealt at once (0: contiguous range per CTA); every thread reads 16 B of each 8 KB row of 512 threads
__global__ void __launch_bounds__(512, 1) k_stream(const uint8_t *w, size_t total, size_t unit, uint32_t *sink) {
    const size_t row = 512 * 16;
    const size_t per_cta = total / gridDim.x;
    uint32_t acc = 0;
    if (unit == 0) {
        const uint8_t *p = w + per_cta * blockIdx.x + threadIdx.x * 16;
        const size_t n = per_cta / row;
        for (size_t i = 0; i + 4 <= n; i += 4) {
            uint4 a = ldg4(p + (i + 0) * row), b = ldg4(p + (i + 1) * row), c = ldg4(p + (i + 2) * row), d = ldg4(p + (i + 3) * row);
            acc += a.x ^ b.y ^ c.z ^ d.w;
        }
    } else {
        const size_t units = total / unit, rows_per_unit = unit / row;
        for (size_t u = blockIdx.x; u < units; u += gridDim.x) {
            const uint8_t *p = w + u * unit + threadIdx.x * 16;
            for (size_t i = 0; i + 4 <= rows_per_unit; i += 4) {
                uint4 a = ldg4(p + (i + 0) * row), b = ldg4(p + (i + 1) * row), c = ldg4(p + (i + 2) * row), d = ldg4(p + (i + 3) * row);
                acc += a.x ^ b.y ^ c.z ^ d.w;
            }
        }
    }
    if (acc == 0x12345678u) sink[0] = acc;
}
int main() {
    const size_t total = (size_t)148 * 64 * 32768;   // 310 MB, a multiple of everything below
    uint8_t *w; uint32_t *sink;
    cudaMalloc(&w, total); cudaMalloc(&sink, 4); cudaMemset(w, 1, total);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const size_t units[] = {0, 32768, 16384, 65536, 131072};
    for (size_t unit : units) {
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            k_stream<<<148, 512>>>(w, total, unit, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep == 2) printf("unit %7zu B: %7.1f us  %6.0f GB/s\n", unit, ms * 1e3, total / (ms * 1e-3) / 1e9);
        }
    }
    return 0;
}
