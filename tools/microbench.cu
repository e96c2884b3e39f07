// microbench.cu — B200 measurements that drive the decode-kernel design (run under gpurun; not part of the library).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/microbench tools/microbench.cu
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <stdio.h>
#include <stdint.h>
#include <vector>
namespace cg = cooperative_groups;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

static float time_ms(cudaStream_t st, void (*fn)(void *), void *arg, int reps) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    fn(arg);
    CK(cudaStreamSynchronize(st));
    CK(cudaEventRecord(a, st));
    for (int i = 0; i < reps; ++i) fn(arg);
    CK(cudaEventRecord(b, st));
    CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    return ms / reps;
}

// ---------------------------------------------------------------- (a) LDG.128 streaming
template <int U>
__global__ void k_bw_ldg(const uint4 *__restrict__ p, size_t n, uint32_t *out) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0;
    for (; i + (U - 1) * stride < n; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(p + i + u * stride));
#pragma unroll
        for (int u = 0; u < U; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) *out = acc;
}

// ---------------------------------------------------------------- (b) cp.async.bulk ring (producer thread + consumer warps)
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
template <int SLOTS, int CHUNK>
__global__ void __launch_bounds__(288) k_bw_bulk(const char *__restrict__ p, size_t bytes_per_cta, uint32_t *out, int touch) {
    extern __shared__ __align__(128) char sm[];
    uint64_t *full = (uint64_t *)(sm + (size_t)SLOTS * CHUNK);
    uint64_t *empty = full + SLOTS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < SLOTS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const char *base = p + (size_t)blockIdx.x * bytes_per_cta;
    const int nchunks = (int)(bytes_per_cta / CHUNK);
    if (warp == 8) {
        if (lane == 0) {
            for (int c = 0; c < nchunks; ++c) {
                int s = c % SLOTS;
                uint32_t ph = (c / SLOTS) & 1;
                if (c >= SLOTS) mbar_wait(&empty[s], ph ^ 1);
                mbar_expect_tx(&full[s], CHUNK);
                bulk_g2s(sm + (size_t)s * CHUNK, base + (size_t)c * CHUNK, CHUNK, &full[s]);
            }
        }
    } else {
        uint32_t acc = 0;
        for (int c = 0; c < nchunks; ++c) {
            int s = c % SLOTS;
            uint32_t ph = (c / SLOTS) & 1;
            mbar_wait(&full[s], ph);
            if (touch) {
                const uint4 *q = (const uint4 *)(sm + (size_t)s * CHUNK);
                for (int i = threadIdx.x; i < CHUNK / 16; i += 256) { uint4 v = q[i]; acc ^= v.x ^ v.y ^ v.z ^ v.w; }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        if (acc == 0x12345678u) *out = acc;
    }
}

// ---------------------------------------------------------------- (c,d) ALU pipes
__global__ void k_dp4a(int iters, int *out) {
    int a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, a4 = 4, a5 = 5, a6 = 6, a7 = 7;
    uint32_t x = threadIdx.x * 2654435761u, y = blockIdx.x + 12345u;
    for (int i = 0; i < iters; ++i) {
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a0) : "r"(x), "r"(y));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a1) : "r"(x), "r"(y));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a2) : "r"(x), "r"(y));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a3) : "r"(x), "r"(y));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a4) : "r"(x), "r"(y));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a5) : "r"(x), "r"(y));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a6) : "r"(x), "r"(y));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a7) : "r"(x), "r"(y));
    }
    int s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 0x12345678) *out = s;
}
__global__ void k_lop(int iters, int *out) {
    uint32_t a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, a4 = 4, a5 = 5, a6 = 6, a7 = 7;
    uint32_t m = blockIdx.x | 0x0F0F0F0Fu, o = 0x30303030u ^ threadIdx.x;
    for (int i = 0; i < iters; ++i) {
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a0) : "r"(m), "r"(o));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a1) : "r"(m), "r"(o));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a2) : "r"(m), "r"(o));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a3) : "r"(m), "r"(o));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a4) : "r"(m), "r"(o));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a5) : "r"(m), "r"(o));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a6) : "r"(m), "r"(o));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a7) : "r"(m), "r"(o));
    }
    uint32_t s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 0x12345678u) *out = s;
}
__global__ void k_ffma(int iters, float *out) {
    float a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, a4 = 4, a5 = 5, a6 = 6, a7 = 7;
    float x = 1.0001f + blockIdx.x * 1e-9f, y = 0.5f;
    for (int i = 0; i < iters; ++i) {
        a0 = fmaf(a0, x, y); a1 = fmaf(a1, x, y); a2 = fmaf(a2, x, y); a3 = fmaf(a3, x, y);
        a4 = fmaf(a4, x, y); a5 = fmaf(a5, x, y); a6 = fmaf(a6, x, y); a7 = fmaf(a7, x, y);
    }
    float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 1.2345f) *out = s;
}
__global__ void k_imma(int iters, int *out) {
    int c[4][4] = {};
    uint32_t a0 = threadIdx.x, a1 = 0x01020304u, a2 = 0x7f7f7f7fu, a3 = blockIdx.x, b0 = 0x01010101u, b1 = threadIdx.x * 3u;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    int s = 0;
    for (int j = 0; j < 4; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    if (s == 0x12345678) *out = s;
}

// ---------------------------------------------------------------- (e) grid barrier latency
__global__ void k_gridbar(int n, unsigned *counter, volatile unsigned *gen) {
    for (int i = 0; i < n; ++i) {
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned g = *gen;
            __threadfence();
            if (atomicAdd(counter, 1) == gridDim.x - 1) {
                *counter = 0;
                __threadfence();
                *gen = g + 1;
            } else {
                while (*gen == g) {}
            }
            __threadfence();
        }
        __syncthreads();
    }
}
__global__ void k_gridbar_cg(int n) {
    cg::grid_group g = cg::this_grid();
    for (int i = 0; i < n; ++i) g.sync();
}

// ---------------------------------------------------------------- (f) dependent tiny kernels
__global__ void k_tiny(int *p) { if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1; }
__global__ void k_tiny_pdl(int *p) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    printf("device %s sms %d clock %d kHz\n", prop.name, sms, prop.clockRate);
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    uint32_t *dout;
    CK(cudaMalloc(&dout, 64));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float ms;

    // (a) LDG
    size_t bytes = (size_t)4 << 30;
    char *buf;
    CK(cudaMalloc(&buf, bytes));
    CK(cudaMemset(buf, 1, bytes));
    for (int ctas_per_sm : {1, 2, 4, 8}) {
        for (int threads : {256, 512}) {
            auto run = [&](int U) {
                int grid = sms * ctas_per_sm;
                for (int rep = 0; rep < 2; ++rep) {
                    CK(cudaEventRecord(e0, st));
                    if (U == 4) k_bw_ldg<4><<<grid, threads, 0, st>>>((const uint4 *)buf, bytes / 16, dout);
                    else k_bw_ldg<8><<<grid, threads, 0, st>>>((const uint4 *)buf, bytes / 16, dout);
                    CK(cudaEventRecord(e1, st));
                    CK(cudaEventSynchronize(e1));
                }
                CK(cudaEventElapsedTime(&ms, e0, e1));
                printf("ldg128 U=%d ctas/sm=%d threads=%d : %.1f GB/s\n", U, ctas_per_sm, threads, bytes / ms / 1e6);
            };
            run(4);
            run(8);
        }
    }
    // small-size LDG: how fast does a 9.4 MB / 33 MB / 66 MB read complete (latency-dominated regime)
    for (size_t mb : {9, 33, 66, 132, 525}) {
        size_t b = mb << 20;
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0, st));
            k_bw_ldg<4><<<sms * 4, 512, 0, st>>>((const uint4 *)(buf + ((size_t)rep << 30)), b / 16, dout);
            CK(cudaEventRecord(e1, st));
            CK(cudaEventSynchronize(e1));
        }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("ldg128 single launch %zu MB: %.2f us  -> %.1f GB/s\n", mb, ms * 1e3, b / ms / 1e6);
    }
    // (b) bulk ring
    {
        auto runb = [&](auto kern, int slots, int chunk, int touch) {
            size_t smem = (size_t)slots * chunk + slots * 16;
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            size_t per = (bytes / sms) / chunk * chunk;
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaEventRecord(e0, st));
                kern<<<sms, 288, smem, st>>>(buf, per, dout, touch);
                CK(cudaEventRecord(e1, st));
                CK(cudaEventSynchronize(e1));
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("bulk ring slots=%d chunk=%d touch=%d: %.1f GB/s\n", slots, chunk, touch, per * sms / ms / 1e6);
        };
        // per-SM streaming ceiling: fewer CTAs, same ring (does an SM go faster when its neighbours are idle?)
        for (int grid : {16, 37, 74, 111}) {
            auto kern = k_bw_bulk<12, 16384>;
            size_t smem = (size_t)12 * 16384 + 12 * 16;
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            size_t per = ((size_t)1 << 30) / grid / 16384 * 16384;
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaEventRecord(e0, st));
                kern<<<grid, 288, smem, st>>>(buf, per, dout, 1);
                CK(cudaEventRecord(e1, st));
                CK(cudaEventSynchronize(e1));
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("bulk ring 12x16K touch=1 grid=%d: %.1f GB/s total, %.1f GB/s per SM\n", grid, per * grid / ms / 1e6, per / ms / 1e6);
        }
        runb(k_bw_bulk<4, 16384>, 4, 16384, 0);
        runb(k_bw_bulk<8, 16384>, 8, 16384, 0);
        runb(k_bw_bulk<12, 16384>, 12, 16384, 0);
        runb(k_bw_bulk<12, 16384>, 12, 16384, 1);
        runb(k_bw_bulk<24, 8192>, 24, 8192, 0);
        runb(k_bw_bulk<24, 8192>, 24, 8192, 1);
        runb(k_bw_bulk<48, 4096>, 48, 4096, 1);
        runb(k_bw_bulk<96, 2048>, 96, 2048, 1);
    }
    // (c,d) pipes: ops per clock per SM
    {
        int iters = 4096, threads = 1024;
        auto rate = [&](const char *name, float ms_, double ops_per_thread_iter) {
            double ops = (double)sms * threads * iters * ops_per_thread_iter;
            double clk = ms_ * 1e-3 * prop.clockRate * 1e3;
            printf("%s: %.3f ms, %.1f lane-ops/clk/SM (at nominal %d MHz)\n", name, ms_, ops / clk / sms, prop.clockRate / 1000);
        };
        for (int rep = 0; rep < 2; ++rep) { CK(cudaEventRecord(e0, st)); k_dp4a<<<sms, threads, 0, st>>>(iters, (int *)dout); CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); }
        CK(cudaEventElapsedTime(&ms, e0, e1)); rate("dp4a", ms, 8);
        for (int rep = 0; rep < 2; ++rep) { CK(cudaEventRecord(e0, st)); k_lop<<<sms, threads, 0, st>>>(iters, (int *)dout); CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); }
        CK(cudaEventElapsedTime(&ms, e0, e1)); rate("lop3", ms, 8);
        for (int rep = 0; rep < 2; ++rep) { CK(cudaEventRecord(e0, st)); k_ffma<<<sms, threads, 0, st>>>(iters, (float *)dout); CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); }
        CK(cudaEventElapsedTime(&ms, e0, e1)); rate("ffma", ms, 8);
        for (int rep = 0; rep < 2; ++rep) { CK(cudaEventRecord(e0, st)); k_imma<<<sms, threads, 0, st>>>(iters, (int *)dout); CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        {
            double mmas = (double)sms * (threads / 32) * iters * 4;
            double clk = ms * 1e-3 * prop.clockRate * 1e3;
            printf("imma m16n8k32: %.3f ms, %.3f warp-mma/clk/SM = %.0f int8 MAC/clk/SM\n", ms, mmas / clk / sms, mmas / clk / sms * 16 * 8 * 32);
        }
    }
    // (e) grid barriers
    {
        unsigned *ctr;
        CK(cudaMalloc(&ctr, 8));
        CK(cudaMemset(ctr, 0, 8));
        int n = 2000;
        for (int threads : {128, 288, 1024}) {
            void *args[] = {&n, &ctr, (void *)&ctr};
            unsigned *gen = ctr + 1;
            void *args2[] = {&n, &ctr, &gen};
            (void)args;
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaEventRecord(e0, st));
                CK(cudaLaunchCooperativeKernel((void *)k_gridbar, dim3(sms), dim3(threads), args2, 0, st));
                CK(cudaEventRecord(e1, st));
                CK(cudaEventSynchronize(e1));
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("grid barrier (atomic+spin) %d CTAs x %d thr: %.3f us each\n", sms, threads, ms * 1e3 / n);
            void *args3[] = {&n};
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaEventRecord(e0, st));
                CK(cudaLaunchCooperativeKernel((void *)k_gridbar_cg, dim3(sms), dim3(threads), args3, 0, st));
                CK(cudaEventRecord(e1, st));
                CK(cudaEventSynchronize(e1));
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("grid barrier (cg grid.sync) %d CTAs x %d thr: %.3f us each\n", sms, threads, ms * 1e3 / n);
        }
    }
    // (f) dependent tiny kernels: stream, graph, graph+PDL
    {
        int *p;
        CK(cudaMalloc(&p, 4));
        CK(cudaMemset(p, 0, 4));
        int n = 400;
        for (int rep = 0; rep < 2; ++rep) {
            CK(cudaEventRecord(e0, st));
            for (int i = 0; i < n; ++i) k_tiny<<<148, 128, 0, st>>>(p);
            CK(cudaEventRecord(e1, st));
            CK(cudaEventSynchronize(e1));
        }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("stream launches: %.3f us per dependent tiny kernel\n", ms * 1e3 / n);
        cudaGraph_t g; cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < n; ++i) k_tiny<<<148, 128, 0, st>>>(p);
        CK(cudaStreamEndCapture(st, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0, st));
            CK(cudaGraphLaunch(ge, st));
            CK(cudaEventRecord(e1, st));
            CK(cudaEventSynchronize(e1));
        }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("graph: %.3f us per dependent tiny kernel\n", ms * 1e3 / n);
        // PDL in a graph
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < n; ++i) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(148); cfg.blockDim = dim3(128); cfg.stream = st;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            CK(cudaLaunchKernelEx(&cfg, k_tiny_pdl, p));
        }
        cudaGraph_t g2; cudaGraphExec_t ge2;
        CK(cudaStreamEndCapture(st, &g2));
        CK(cudaGraphInstantiate(&ge2, g2, 0));
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0, st));
            CK(cudaGraphLaunch(ge2, st));
            CK(cudaEventRecord(e1, st));
            CK(cudaEventSynchronize(e1));
        }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        int hv; CK(cudaMemcpy(&hv, p, 4, cudaMemcpyDeviceToHost));
        printf("graph + PDL: %.3f us per dependent tiny kernel (counter %d)\n", ms * 1e3 / n, hv);
    }
    printf("done\n");
    return 0;
}
