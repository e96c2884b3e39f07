// microbench2.cu — what bounds the GEMV consumer loop?  Runs the block-group loop of csrc/gemv.cu on a resident shared-memory
// chunk (no streaming) with 8/16/24 warps and with pieces removed.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdint.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
#define MMA_MAGIC 0x4B400000
#define MMA_MAGIC_F 12582912.0f
__device__ __forceinline__ void mma_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// MODE 0: full; 1: no IMMA (xor instead); 2: no epilogue; 3: loads only; 4: no LDS (regs only) but IMMA+epilogue
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k_loop(int iters, float *out, long long *cyc) {
    extern __shared__ __align__(128) uint8_t sm[];
    uint8_t *slot = sm;                       // 18432 B chunk
    uint8_t *sact = sm + 18432;               // 16 bgs * 384
    float *sscale = (float *)(sact + 16 * 384);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int nw = blockDim.x >> 5;
    for (int i = tid; i < (18432 + 16 * 384 + 256) / 4; i += blockDim.x) ((uint32_t *)sm)[i] = i * 2654435761u;
    for (int i = tid; i < 64; i += blockDim.x) sscale[i] = 1e-3f;
    __syncthreads();
    float acc[4] = {0, 0, 0, 0};
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const uint4 *sc4 = (const uint4 *)(slot + 64 * 256);
        for (int bg = warp; bg < 16; bg += nw) {
            const uint4 *nib = (const uint4 *)(slot + (size_t)bg * 1024);
            uint4 wa, wb, sd, b01 = make_uint4(0, 0, 0, 0), b23 = make_uint4(0, 0, 0, 0);
            float4 as4;
            if (MODE != 4) {
                wa = nib[lane]; wb = nib[32 + lane]; sd = sc4[bg * 8 + g]; as4 = *(const float4 *)(sscale + bg * 4);
                if (g < 3) { const uint4 *bp = (const uint4 *)(sact + (size_t)bg * 384 + g * 128 + t * 32); b01 = bp[0]; b23 = bp[1]; }
            } else {
                wa = make_uint4(it, lane, bg, 7); wb = make_uint4(lane, it, 3, bg); sd = make_uint4(0x3c003c00, 0x3c003c00, 0x3c003c00, 0x3c003c00);
                as4 = make_float4(1e-3f, 1e-3f, 1e-3f, 1e-3f); b01 = make_uint4(it, 1, 2, 3); b23 = b01;
            }
            const uint32_t ua[4] = {wa.x, wa.y, wa.z, wa.w}, ub[4] = {wb.x, wb.y, wb.z, wb.w};
            const uint32_t bx[4] = {b01.x, b01.z, b23.x, b23.z}, by[4] = {b01.y, b01.w, b23.y, b23.w};
            const uint32_t sdw[4] = {sd.x, sd.y, sd.z, sd.w};
            const float asf[4] = {as4.x, as4.y, as4.z, as4.w};
            if (MODE == 3) { acc[0] += __uint_as_float(wa.x ^ wb.y ^ sd.z ^ b01.x ^ b23.w) + as4.x; continue; }
            int c[4][4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                c[j][0] = c[j][1] = c[j][2] = c[j][3] = MMA_MAGIC;
                uint32_t a0 = (ua[j] << 4) & 0xF0F0F0F0u, a1 = (ub[j] << 4) & 0xF0F0F0F0u, a2 = ua[j] & 0xF0F0F0F0u, a3 = ub[j] & 0xF0F0F0F0u;
                if (MODE == 1) { c[j][0] += (a0 ^ bx[j]) & 0xFFFF; c[j][1] += (a1 ^ by[j]) & 0xFFFF; c[j][2] += a2 & 0xFFFF; c[j][3] += a3 & 0xFFFF; }
                else mma_s8(c[j], a0, a1, a2, a3, bx[j], by[j]);
            }
            if (MODE == 2) { acc[0] += __int_as_float(c[0][0] ^ c[1][1] ^ c[2][2] ^ c[3][3]); continue; }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 dd = __half22float2(*(const __half2 *)&sdw[j]);
                const float sb = asf[j] * 0.0625f;
                const float f0 = dd.x * sb, f1 = dd.y * sb;
                acc[0] = fmaf(__int_as_float(c[j][0]) - MMA_MAGIC_F, f0, acc[0]);
                acc[1] = fmaf(__int_as_float(c[j][1]) - MMA_MAGIC_F, f0, acc[1]);
                acc[2] = fmaf(__int_as_float(c[j][2]) - MMA_MAGIC_F, f1, acc[2]);
                acc[3] = fmaf(__int_as_float(c[j][3]) - MMA_MAGIC_F, f1, acc[3]);
            }
        }
        __syncwarp();
    }
    long long t1 = clock64();
    if (tid == 0) cyc[blockIdx.x] = t1 - t0;
    if (acc[0] + acc[1] + acc[2] + acc[3] == 1.2345f) *out = acc[0];
}
template <int MODE>
void run(const char *name, int warps, float *dout, long long *dcyc) {
    int iters = 2000;
    size_t smem = 18432 + 16 * 384 + 256 + 128;
    CK(cudaFuncSetAttribute(k_loop<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_loop<MODE><<<148, warps * 32, smem>>>(iters, dout, dcyc);
    CK(cudaDeviceSynchronize());
    long long h[148];
    CK(cudaMemcpy(h, dcyc, sizeof h, cudaMemcpyDeviceToHost));
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += h[i];
    printf("%-28s warps=%2d : %.0f cycles per 64-block chunk\n", name, warps, avg / 148 / iters);
}
int main() {
    float *dout; long long *dcyc;
    CK(cudaMalloc(&dout, 64)); CK(cudaMalloc(&dcyc, 148 * 8));
    for (int w : {4, 8, 16, 32}) {
        run<0>("full", w, dout, dcyc);
        run<1>("no IMMA", w, dout, dcyc);
        run<2>("no epilogue", w, dout, dcyc);
        run<3>("loads only", w, dout, dcyc);
        run<4>("no LDS", w, dout, dcyc);
    }
    return 0;
}
