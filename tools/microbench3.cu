// microbench3.cu — streaming GEMV consumer structures, end to end from HBM (2 GB buffer, 148 persistent CTAs):
//   ldg<D>   : every warp streams its block-group column straight into registers (LDG.128, D block-groups in flight),
//              activation fragments resident in registers, no shared-memory ring
//   ring     : cp.async.bulk ring (as csrc/gemv.cu) with activation fragments in registers (3 LDS.128 per block-group)
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdint.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
#define MMA_MAGIC 0x4B400000
#define MMA_MAGIC_F 12582912.0f
#define CHUNK 18432
__device__ __forceinline__ void mma_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
struct W { uint4 wa, wb, sd; };
__device__ __forceinline__ void compute(const W &r, const uint4 &b01, const uint4 &b23, const float4 &as4, float (&acc)[4]) {
    const uint32_t ua[4] = {r.wa.x, r.wa.y, r.wa.z, r.wa.w}, ub[4] = {r.wb.x, r.wb.y, r.wb.z, r.wb.w};
    const uint32_t bx[4] = {b01.x, b01.z, b23.x, b23.z}, by[4] = {b01.y, b01.w, b23.y, b23.w};
    const uint32_t sdw[4] = {r.sd.x, r.sd.y, r.sd.z, r.sd.w};
    const float asf[4] = {as4.x, as4.y, as4.z, as4.w};
    int c[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        c[j][0] = c[j][1] = c[j][2] = c[j][3] = MMA_MAGIC;
        mma_s8(c[j], (ua[j] << 4) & 0xF0F0F0F0u, (ub[j] << 4) & 0xF0F0F0F0u, ua[j] & 0xF0F0F0F0u, ub[j] & 0xF0F0F0F0u, bx[j], by[j]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 dd = __half22float2(*(const __half2 *)&sdw[j]);
        const float f0 = dd.x * asf[j], f1 = dd.y * asf[j];
        acc[0] = fmaf(__int_as_float(c[j][0]) - MMA_MAGIC_F, f0, acc[0]);
        acc[1] = fmaf(__int_as_float(c[j][1]) - MMA_MAGIC_F, f0, acc[1]);
        acc[2] = fmaf(__int_as_float(c[j][2]) - MMA_MAGIC_F, f1, acc[2]);
        acc[3] = fmaf(__int_as_float(c[j][3]) - MMA_MAGIC_F, f1, acc[3]);
    }
}
// chunk layout as csrc/layout.cuh: 16 block-groups x 1024 B nibbles, then 16 x 128 B scales
template <int D>
__global__ void __launch_bounds__(512, 1) k_ldg(const uint8_t *__restrict__ w, int chunks_per_cta, float *out) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2;
    const uint8_t *base = w + (size_t)blockIdx.x * chunks_per_cta * CHUNK;
    uint4 b01[2], b23[2];
    float4 as4[2];
    for (int i = 0; i < 2; ++i) { b01[i] = make_uint4(lane, i, 3, 4); b23[i] = make_uint4(i, lane, 1, 2); as4[i] = make_float4(1e-3f, 2e-3f, 1e-3f, 3e-3f); }
    float acc[4] = {0, 0, 0, 0};
    W buf[D];
    auto load = [&](int c, W &r) {
        const uint8_t *ch = base + (size_t)c * CHUNK;
        const uint4 *nib = (const uint4 *)(ch + warp * 1024);
        r.wa = ldg_stream(nib + lane);
        r.wb = ldg_stream(nib + 32 + lane);
        r.sd = ldg_stream((const uint4 *)(ch + 16384) + warp * 8 + g);
    };
#pragma unroll
    for (int d = 0; d < D; ++d) load(d, buf[d]);
    for (int c = 0; c < chunks_per_cta; c += D) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            compute(buf[d], b01[(c + d) & 1], b23[(c + d) & 1], as4[(c + d) & 1], acc);
            if (c + d + D < chunks_per_cta) load(c + d + D, buf[d]);
        }
    }
    if (acc[0] + acc[1] + acc[2] + acc[3] == 1.2345f) *out = acc[0];
}
// ---- ring variant
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t *b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
template <int S>
__global__ void __launch_bounds__(544, 1) k_ring(const uint8_t *__restrict__ w, int chunks_per_cta, float *out) {
    extern __shared__ __align__(128) uint8_t sm[];
    uint64_t *full = (uint64_t *)(sm + (size_t)S * CHUNK), *empty = full + S;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2;
    if (tid == 0) { for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 16); } asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    const uint8_t *base = w + (size_t)blockIdx.x * chunks_per_cta * CHUNK;
    if (warp == 16) {
        if (lane == 0)
            for (int c = 0; c < chunks_per_cta; ++c) {
                int s = c % S;
                if (c >= S) mbar_wait(&empty[s], ((c / S) & 1) ^ 1);
                mbar_expect_tx(&full[s], CHUNK);
                bulk_g2s(sm + (size_t)s * CHUNK, base + (size_t)c * CHUNK, CHUNK, &full[s]);
            }
        return;
    }
    uint4 b01[2], b23[2];
    float4 as4[2];
    for (int i = 0; i < 2; ++i) { b01[i] = make_uint4(lane, i, 3, 4); b23[i] = make_uint4(i, lane, 1, 2); as4[i] = make_float4(1e-3f, 2e-3f, 1e-3f, 3e-3f); }
    float acc[4] = {0, 0, 0, 0};
    for (int c = 0; c < chunks_per_cta; ++c) {
        int s = c % S;
        mbar_wait(&full[s], (c / S) & 1);
        const uint8_t *ch = sm + (size_t)s * CHUNK;
        W r;
        const uint4 *nib = (const uint4 *)(ch + warp * 1024);
        r.wa = nib[lane]; r.wb = nib[32 + lane]; r.sd = ((const uint4 *)(ch + 16384))[warp * 8 + g];
        compute(r, b01[c & 1], b23[c & 1], as4[c & 1], acc);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
    if (acc[0] + acc[1] + acc[2] + acc[3] == 1.2345f) *out = acc[0];
}
int main() {
    size_t bytes = (size_t)2 << 30;
    uint8_t *buf; float *dout;
    CK(cudaMalloc(&buf, bytes)); CK(cudaMemset(buf, 0x11, bytes)); CK(cudaMalloc(&dout, 64));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    int cpc = (int)(bytes / 148 / CHUNK);
    cpc -= cpc % 12;
    float ms;
    auto report = [&](const char *name) { CK(cudaEventElapsedTime(&ms, e0, e1)); printf("%-22s %.1f GB/s (%.0f cycles/chunk at 1.965 GHz)\n", name, (double)cpc * 148 * CHUNK / ms / 1e6, ms * 1e-3 * 1.965e9 / cpc); };
#define RUN(name, launch) for (int rep = 0; rep < 2; ++rep) { CK(cudaEventRecord(e0)); launch; CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); } report(name);
    RUN("ldg D=2", (k_ldg<2><<<148, 512>>>(buf, cpc, dout)));
    RUN("ldg D=3", (k_ldg<3><<<148, 512>>>(buf, cpc, dout)));
    RUN("ldg D=4", (k_ldg<4><<<148, 512>>>(buf, cpc, dout)));
    RUN("ldg D=6", (k_ldg<6><<<148, 512>>>(buf, cpc, dout)));
    { size_t smem = 10 * CHUNK + 256; CK(cudaFuncSetAttribute(k_ring<10>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      RUN("ring S=10 (B in regs)", (k_ring<10><<<148, 544, smem>>>(buf, cpc, dout))); }
    { size_t smem = 6 * CHUNK + 256; CK(cudaFuncSetAttribute(k_ring<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      RUN("ring S=6 (B in regs)", (k_ring<6><<<148, 544, smem>>>(buf, cpc, dout))); }
    return 0;
}
