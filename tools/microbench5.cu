// microbench5.cu — what does an all-to-all hand-over of a 4096-float vector between 148 RESIDENT CTAs cost on one B200?
// (VERDICT r1 #4c: the price of a stage boundary inside a persistent decode kernel, against ~2 us of a programmatic-dependent-
// launch boundary and 2-3 us measured for cooperative_groups grid.sync() under load.)
// Every epoch each CTA writes its 1/148 of the vector, then needs the WHOLE vector (as a GEMV needs its whole input):
//   V0  plain stores, cg::grid.sync(), ld.cg of the vector
//   V1  tagged pairs (value, epoch) stored with st.volatile.v2, no fence, no barrier: readers re-read until every tag matches
//   V2  plain stores, __syncthreads, thread 0: __threadfence + flag[cta] = epoch; warp 0 of every CTA polls the 148 flags,
//       __syncthreads, ld.cg of the vector
// Each variant runs alone and with a weight stream around it (every thread requests 4 x 16 B of HBM per epoch BEFORE the
// hand-over and consumes them after it, as the GEMV's register pipeline does): 32 KB per CTA and epoch.
// Two vector buffers alternate, which is enough: nobody can be two epochs ahead of anybody else.
// Spins are bounded (an error word is set instead of hanging the GPU).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/microbench5.cu -o tools/bin/microbench5
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;
#define N 4096
#define SPIN_MAX 20000000u
__device__ __forceinline__ uint4 ldg4(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint4 ldv4(const void *p) {
    uint4 r;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void stv2(void *p, uint32_t x, uint32_t y) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ uint32_t ldv(const void *p) {
    uint32_t r;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
    return r;
}

template <int V, bool STREAM>
__global__ void __launch_bounds__(512, 1) k_handover(float *vec, uint2 *pairs, uint32_t *flags, const uint8_t *w, size_t w_per_cta, int epochs,
                                                     uint32_t *err, float *sink) {
    __shared__ float sx[N];
    const int tid = threadIdx.x, bid = blockIdx.x, nb = gridDim.x;
    const int r0 = (int)((long long)N * bid / nb), r1 = (int)((long long)N * (bid + 1) / nb);
    float acc = 0.f;
    uint32_t wacc = 0;
    const uint8_t *wp = w + w_per_cta * bid + tid * 16;
    size_t woff = 0;
    for (int e = 1; e <= epochs; ++e) {
        uint4 q0, q1, q2, q3;
        if (STREAM) {   // next piece of the weight stream: requested before the hand-over, consumed after it
            q0 = ldg4(wp + woff); q1 = ldg4(wp + woff + 8192); q2 = ldg4(wp + woff + 16384); q3 = ldg4(wp + woff + 24576);
            woff += 32768;
            if (woff + 32768 > w_per_cta) woff = 0;
        }
        float *vb = vec + (e & 1) * N;
        uint2 *pb = pairs + (e & 1) * N;
        // ---- produce my rows (value depends on the previous epoch's sum: a real dependency chain)
        if (V == 1) {
            for (int i = r0 + tid; i < r1; i += 512) stv2(pb + i, __float_as_uint(acc * 1e-9f + (float)i), (uint32_t)e);
        } else {
            for (int i = r0 + tid; i < r1; i += 512) __stcg(vb + i, acc * 1e-9f + (float)i);
        }
        // ---- hand-over
        if (V == 0) {
            cg::this_grid().sync();
        } else if (V == 2) {
            __syncthreads();
            if (tid == 0) { __threadfence(); asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(flags + bid), "r"((uint32_t)e) : "memory"); }
            if (tid < 32) {
                uint32_t spins = 0;
                for (;;) {
                    bool ok = true;
                    for (int f = tid; f < nb; f += 32) ok = ok && ldv(flags + f) >= (uint32_t)e;
                    if (__all_sync(0xffffffffu, ok)) break;
                    if (++spins > SPIN_MAX) { *err = 2; break; }
                }
                __threadfence();
            }
            __syncthreads();
        }
        // ---- consume the whole vector
        if (V == 1) {
            for (int q = tid; q < N / 2; q += 512) {
                uint4 v;
                uint32_t spins = 0;
                for (;;) {
                    v = ldv4((const uint4 *)pb + q);
                    if (v.y == (uint32_t)e && v.w == (uint32_t)e) break;
                    if (++spins > SPIN_MAX) { *err = 1; break; }
                }
                sx[2 * q] = __uint_as_float(v.x);
                sx[2 * q + 1] = __uint_as_float(v.z);
            }
        } else {
            for (int q = tid; q < N / 4; q += 512) ((float4 *)sx)[q] = __ldcg((const float4 *)vb + q);
        }
        __syncthreads();
        float s = 0.f;
        for (int i = tid & 31; i < N; i += 32 * 16) s += sx[i];   // a little dependent work on the vector
        acc += s;
        if (STREAM) wacc += q0.x ^ q1.y ^ q2.z ^ q3.w;
        __syncthreads();
    }
    if (acc == 1.2345f || wacc == 0x12345678u) sink[0] = acc + (float)wacc;
}

template <int V, bool STREAM>
static void run(const char *name, float *vec, uint2 *pairs, uint32_t *flags, const uint8_t *w, size_t w_per_cta, uint32_t *err, float *sink, int grid) {
    int epochs = 2000;
    cudaMemset(pairs, 0, 2 * N * 8);
    cudaMemset(flags, 0, 1024 * 4);
    cudaMemset(err, 0, 4);
    void *args[] = {&vec, &pairs, &flags, &w, &w_per_cta, &epochs, &err, &sink};
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaMemset(pairs, 0, 2 * N * 8);
        cudaMemset(flags, 0, 1024 * 4);
        cudaEventRecord(e0);
        cudaError_t le = cudaLaunchCooperativeKernel((const void *)k_handover<V, STREAM>, dim3(grid), dim3(512), args, 0, 0);
        cudaEventRecord(e1);
        cudaError_t se = cudaEventSynchronize(e1);
        if (le != cudaSuccess || se != cudaSuccess) { printf("%s: launch %s / sync %s\n", name, cudaGetErrorString(le), cudaGetErrorString(se)); return; }
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    uint32_t herr = 0;
    cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost);
    printf("%-44s %7.3f us per epoch%s%s\n", name, best * 1e3 / epochs, STREAM ? "  (32 KB of HBM per CTA and epoch: 4.85 MB = 0.74 us at 6.55 TB/s)" : "",
           herr ? "  [SPIN LIMIT HIT]" : "");
}

int main() {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = sms;
    const size_t w_per_cta = (size_t)8 << 20;   // 8 MB per CTA: 1.2 GB in all, far beyond L2
    float *vec, *sink; uint2 *pairs; uint32_t *flags, *err; uint8_t *w;
    cudaMalloc(&vec, 2 * N * 4); cudaMalloc(&pairs, 2 * N * 8); cudaMalloc(&flags, 1024 * 4); cudaMalloc(&err, 4); cudaMalloc(&sink, 4);
    cudaMalloc(&w, w_per_cta * grid);
    cudaMemset(w, 1, w_per_cta * grid);
    cudaMemset(vec, 0, 2 * N * 4);
    printf("hand-over of a %d-float vector between %d resident CTAs of 512 threads\n", N, grid);
    run<0, false>("V0 grid.sync", vec, pairs, flags, w, w_per_cta, err, sink, grid);
    run<1, false>("V1 tagged pairs", vec, pairs, flags, w, w_per_cta, err, sink, grid);
    run<2, false>("V2 fence + per-CTA flags", vec, pairs, flags, w, w_per_cta, err, sink, grid);
    run<0, true>("V0 grid.sync            + weight stream", vec, pairs, flags, w, w_per_cta, err, sink, grid);
    run<1, true>("V1 tagged pairs         + weight stream", vec, pairs, flags, w, w_per_cta, err, sink, grid);
    run<2, true>("V2 fence + per-CTA flags + weight stream", vec, pairs, flags, w, w_per_cta, err, sink, grid);
    return 0;
}
