// gemv.cu — the streaming Q4_0 GEMV of the decode path (stages a3.3 QKV, a3.6 O, a3.7 gate/up+act*mul,
// a3.8 down and LM head of Ollama's /api/generate loop; SURVEY.md §3.3, call site experiment/RunnerConfig.py:128-131).
//
// One persistent CTA per SM.  Warp 8 lane 0 is the producer: it streams the CTA's contiguous range of
// weight tiles from HBM into a shared-memory ring with cp.async.bulk (UBLKCP), one 18 KB chunk per copy,
// completion on an mbarrier.  Weights do not depend on the previous kernel, so the producer starts before
// griddepcontrol.wait: under programmatic dependent launch the ring fills while the previous kernel drains.
// Warps 0-7 are consumers: per 32-wide block they rebuild int8 A fragments from the nibbles (two's-complement
// nibble in the high half of each byte = 16*(q-8): 3 ALU ops per 8 weights), load the activation digits as the
// B fragment and issue mma.sync.m16n8k32.s8.s8.s32: 16 rows x 32 weights x 3 digit planes of exact integer
// dot products per instruction.  The int32 results are scaled by d_w * 2^e / 16 into fp32 accumulators; the
// eight warps split the K dimension and are reduced through shared memory at the end of each 16-row tile.
#include "kernels.cuh"
#include "layout.cuh"
#include "randinit.cuh"

// ------------------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ void mma_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// ------------------------------------------------------------------------------------------ pack / init
// GGML Q4_0 row-major blocks (18 B: f16 d | 16 B nibbles) -> streaming layout (layout.cuh). One thread per block.
__global__ void k_pack_q4_0(const uint8_t *__restrict__ src, int64_t rows, int64_t nb, int64_t nbp, uint8_t *__restrict__ dst,
                            int64_t dst_row0, int64_t dst_row_stride) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    int64_t r = i / nb, b = i - r * nb;
    const uint8_t *p = src + i * 18;
    PackedLoc loc = packed_loc(dst_row0 + r * dst_row_stride, b, nbp);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        uint32_t w = (uint32_t)p[2 + 4 * t] | ((uint32_t)p[3 + 4 * t] << 8) | ((uint32_t)p[4 + 4 * t] << 16) | ((uint32_t)p[5 + 4 * t] << 24);
        *(uint32_t *)(dst + loc.word0 + t * 16) = w ^ 0x88888888u;
    }
    *(uint16_t *)(dst + loc.scale) = (uint16_t)p[0] | ((uint16_t)p[1] << 8);
}
void launch_pack_q4_0(const void *d_src, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st) {
    int64_t nb = cols / QK, n = rows * nb;
    k_pack_q4_0<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const uint8_t *)d_src, rows, nb, nblocks_padded(cols), (uint8_t *)d_packed, dst_row0, dst_row_stride);
}
// Deterministic random init straight into the streaming layout; same bytes as b2_randinit_tensor + pack.
__global__ void k_randinit_q4_0(uint64_t key, float sigma, int64_t rows, int64_t nb, int64_t nbp, uint8_t *__restrict__ dst,
                                int64_t dst_row0, int64_t dst_row_stride) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    int64_t r = i / nb, b = i - r * nb;
    float x[QK];
    ri_gauss_block(key, (uint64_t)i * QK, sigma, x);
    uint16_t dbits;
    uint32_t w[4];
    ri_quant_q4_0(x, dbits, w);
    PackedLoc loc = packed_loc(dst_row0 + r * dst_row_stride, b, nbp);
#pragma unroll
    for (int t = 0; t < 4; ++t) *(uint32_t *)(dst + loc.word0 + t * 16) = w[t] ^ 0x88888888u;
    *(uint16_t *)(dst + loc.scale) = dbits;
}
void launch_randinit_q4_0(uint64_t key, float sigma, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st) {
    int64_t nb = cols / QK, n = rows * nb;
    k_randinit_q4_0<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(key, sigma, rows, nb, nblocks_padded(cols), (uint8_t *)d_packed, dst_row0, dst_row_stride);
}
__global__ void k_randinit_f32(uint64_t key, float sigma, float offset, int64_t n, float *__restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ri_add(ri_mul(ri_gauss(key, (uint64_t)i), sigma), offset);
}
void launch_randinit_f32(uint64_t key, float sigma, float offset, int64_t n, float *d_out, cudaStream_t st) {
    k_randinit_f32<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(key, sigma, offset, n, d_out);
}
// one row of a packed matrix, dequantised (token embedding lookup and tests)
__device__ __forceinline__ float packed_elem(const uint8_t *w, int64_t row, int64_t col, int64_t nbp) {
    PackedLoc loc = packed_loc(row, col / QK, nbp);
    int e = (int)(col % QK);
    uint32_t word = *(const uint32_t *)(w + loc.word0 + ((e & 15) >> 2) * 16);
    int nib = (word >> (8 * (e & 3) + ((e >> 4) << 2))) & 0xF;    // two's complement of q-8
    int v = (nib ^ 8) - 8;
    return (float)v * __half2float(__ushort_as_half(*(const uint16_t *)(w + loc.scale)));
}

// ------------------------------------------------------------------------------------------ the kernel
#define GEMV_MAX_SLOTS 6
#define GEMV_THREADS 288
#define MMA_MAGIC 0x4B400000          // int bits of 12582912.0f: accumulators start here, so c reinterpreted as float is 12582912 + S exactly
#define MMA_MAGIC_F 12582912.0f

struct GemvArgs {
    const uint8_t *w;        // packed matrix
    int rows, nbp, n_tiles;  // logical rows, padded blocks per row, tiles
    int tile_align;          // tiles are dealt to CTAs in multiples of this (4 for act*mul so A3 blocks stay whole)
    int n_slots;             // ring depth (chunks of CHUNK_BYTES)
    // input
    const float *x;          // IN_NORM: residual stream, n_in floats
    const float *norm_w;
    float eps;
    int plus_one, n_in;
    const uint8_t *act_in;   // IN_ACT: A3 vector in global memory (staged into shared memory by the producer)
    // output
    float *y;                // fp32 output (OUT_ACTMUL: optional copy of h)
    const float *bias_or_res;
    uint8_t *act_out;        // OUT_ACTMUL: A3 of h
    int n_out_act;           // elements of the A3 output vector (for its scale area offset)
    int act;
    float *pmax;             // OUT_LOGITS: per-CTA arg-max partials
    int *pidx;
};
enum { IN_NORM = 0, IN_ACT = 1 };
enum { OUT_BIAS = 0, OUT_RES = 1, OUT_ACTMUL = 2, OUT_LOGITS = 3 };

__device__ __forceinline__ float act_fn(float g, int act) {
    if (act == B2_ACT_SILU) return g / (1.0f + expf(-g));
    float g3 = 0.044715f * g * g * g;
    return 0.5f * g * (1.0f + tanhf(0.7978845608028654f * (g + g3)));
}

// shared memory map (bytes): [ring n_slots*CHUNK_BYTES][red 4096][hbuf 192][barriers 128][A3 of the input nbp*100]
template <int IN, int OUT>
__global__ void __launch_bounds__(GEMV_THREADS, 2) k_gemv(const GemvArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *ring = smem;
    float *red = (float *)(smem + (size_t)a.n_slots * CHUNK_BYTES);            // [2][8 warps][16 rows][4]
    float *hbuf = red + 2 * 8 * 16 * 4;                                         // 32 (+ 8 block-reduce scratch)
    uint64_t *full = (uint64_t *)(hbuf + 48);
    uint64_t *empty = full + GEMV_MAX_SLOTS;
    uint64_t *actbar = empty + GEMV_MAX_SLOTS;
    uint8_t *sact = (uint8_t *)(full + 16);                                     // A3 image of the input vector
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int S = a.n_slots;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 8); }
        mbar_init(actbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        pdl_launch_dependents();
    }
    __syncthreads();

    // contiguous, aligned share of the tiles
    const int units = (a.n_tiles + a.tile_align - 1) / a.tile_align;
    const int u0 = (int)((int64_t)units * blockIdx.x / gridDim.x), u1 = (int)((int64_t)units * (blockIdx.x + 1) / gridDim.x);
    const int t0 = u0 * a.tile_align, t1 = min(a.n_tiles, u1 * a.tile_align);
    const int cpt = (a.nbp + KC - 1) / KC;                                      // chunks per tile
    const int n_chunks = (t1 - t0) * cpt;

    if (warp == 8) {
        // ------------------------------------------------------------------ producer
        if (lane == 0) {
            const uint8_t *base = a.w + (uint64_t)t0 * a.nbp * BLOCK_BYTES;
            auto issue = [&](int c) {
                const int s = c % S;
                const int ck = c % cpt, tile = c / cpt;
                const int cb = min(KC, a.nbp - ck * KC);
                if (c >= S) mbar_wait(&empty[s], ((c / S) & 1) ^ 1);
                mbar_expect_tx(&full[s], (uint32_t)cb * BLOCK_BYTES);
                bulk_g2s(ring + (size_t)s * CHUNK_BYTES, base + ((uint64_t)tile * a.nbp + (uint64_t)ck * KC) * BLOCK_BYTES,
                         (uint32_t)cb * BLOCK_BYTES, &full[s]);
            };
            // weights are immutable: fill the ring before waiting for the previous kernel
            const int pre = min(n_chunks, S);
            for (int c = 0; c < pre; ++c) issue(c);
            if (IN == IN_ACT) {
                pdl_wait();   // the activation vector is the previous kernel's output
                const uint32_t bytes = (uint32_t)a.nbp * 100;
                mbar_expect_tx(actbar, bytes);
                bulk_g2s(sact, a.act_in, bytes, actbar);
            }
            for (int c = pre; c < n_chunks; ++c) issue(c);
        }
        return;
    }

    // ---------------------------------------------------------------------- consumers
    // Vectors written by the previous kernel are read with ld.global.cg (L2): under programmatic dependent launch
    // this CTA may share its SM's L1 with a still-running CTA of that kernel, so L1 must not be trusted for them.
    float *sscale = (float *)(sact + (size_t)a.nbp * 96);
    float sfold;   // 1/16 of the nibble trick, folded into the activation scale
    if (IN == IN_NORM) {
        pdl_wait();   // x is the previous kernel's output
        // RMSNorm (oracle/model.py:_norm op order) + activation grid, redundantly per CTA, into shared memory
        float ss = 0.f;
        for (int i = tid; i < a.n_in; i += 256) { float v = __ldcg(a.x + i); ss = fmaf(v, v, ss); }
        ss = warp_sum(ss);
        if (lane == 0) hbuf[32 + warp] = ss;
        consumer_sync();
        float tot = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < 8; ++w8) tot += hbuf[32 + w8];
        const float r = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(tot, (float)a.n_in), a.eps)));
        const int nb_in = a.n_in / QK;
        for (int b = warp; b < a.nbp; b += 8) {
            float yv = 0.f;
            if (b < nb_in) {
                float wi = a.norm_w[b * QK + lane];
                if (a.plus_one) wi = __fadd_rn(1.0f, wi);
                yv = __fmul_rn(__fmul_rn(__ldcg(a.x + b * QK + lane), r), wi);
            }
            float sc;
            int X = act_grid_lane(yv, sc);
            int d0, d1, d2;
            act_digits(X, d0, d1, d2);
            uint8_t *dst = sact + a3_off(b, 0, lane);
            dst[0] = (uint8_t)d0;
            dst[128] = (uint8_t)d1;
            dst[256] = (uint8_t)d2;
            if (lane == 0) sscale[b] = (b < nb_in) ? sc * 0.0625f : 0.f;
        }
        consumer_sync();
        sfold = 1.0f;
    } else {
        mbar_wait(actbar, 0);
        sfold = 0.0625f;
    }

    float best_v = -INFINITY;
    int best_i = 0x7fffffff;
    int chunk = 0;
    for (int tile = t0; tile < t1; ++tile) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int ck = 0; ck < cpt; ++ck, ++chunk) {
            const int s = chunk % S;
            const int cb = min(KC, a.nbp - ck * KC);
            mbar_wait(&full[s], (chunk / S) & 1);
            const uint8_t *slot = ring + (size_t)s * CHUNK_BYTES;
            const uint4 *sc4 = (const uint4 *)(slot + (size_t)cb * 256);
            for (int bg = warp; bg < (cb >> 2); bg += 8) {
                const int kbg = ck * (KC / 4) + bg;                              // block-group index along K
                // ---- all loads of this block-group first (6 x LDS.128)
                const uint4 *nib = (const uint4 *)(slot + (size_t)bg * 1024);
                const uint4 wa = nib[lane], wb = nib[32 + lane];
                const uint4 sd = sc4[bg * 8 + g];                                // half2 (row g, row g+8) x 4 blocks
                const float4 as4 = *(const float4 *)(sscale + kbg * 4);
                uint4 b01 = make_uint4(0u, 0u, 0u, 0u), b23 = make_uint4(0u, 0u, 0u, 0u);
                if (g < 3) {
                    const uint4 *bp = (const uint4 *)(sact + (size_t)kbg * 384 + g * 128 + t * 32);
                    b01 = bp[0];
                    b23 = bp[1];
                }
                const uint32_t ua[4] = {wa.x, wa.y, wa.z, wa.w}, ub[4] = {wb.x, wb.y, wb.z, wb.w};
                const uint32_t bx[4] = {b01.x, b01.z, b23.x, b23.z}, by[4] = {b01.y, b01.w, b23.y, b23.w};
                const uint32_t sdw[4] = {sd.x, sd.y, sd.z, sd.w};
                const float asf[4] = {as4.x, as4.y, as4.z, as4.w};
                // ---- four independent MMAs
                int c[4][4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    c[j][0] = c[j][1] = c[j][2] = c[j][3] = MMA_MAGIC;
                    mma_s8(c[j], (ua[j] << 4) & 0xF0F0F0F0u, (ub[j] << 4) & 0xF0F0F0F0u, ua[j] & 0xF0F0F0F0u, ub[j] & 0xF0F0F0F0u, bx[j], by[j]);
                }
                // ---- scale into the fp32 accumulators
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 dd = __half22float2(*(const __half2 *)&sdw[j]);
                    const float sb = asf[j] * sfold;
                    const float f0 = dd.x * sb, f1 = dd.y * sb;
                    acc[0] = fmaf(__int_as_float(c[j][0]) - MMA_MAGIC_F, f0, acc[0]);
                    acc[1] = fmaf(__int_as_float(c[j][1]) - MMA_MAGIC_F, f0, acc[1]);
                    acc[2] = fmaf(__int_as_float(c[j][2]) - MMA_MAGIC_F, f1, acc[2]);
                    acc[3] = fmaf(__int_as_float(c[j][3]) - MMA_MAGIC_F, f1, acc[3]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        // ---- cross-warp reduction of the tile (double-buffered scratch: one barrier per tile)
        float *rb = red + (size_t)(tile & 1) * (8 * 16 * 4);
        if (t < 2) {
            rb[(warp * 16 + g) * 4 + 2 * t] = acc[0];
            rb[(warp * 16 + g) * 4 + 2 * t + 1] = acc[1];
            rb[(warp * 16 + g + 8) * 4 + 2 * t] = acc[2];
            rb[(warp * 16 + g + 8) * 4 + 2 * t + 1] = acc[3];
        }
        consumer_sync();
        if (warp == 0) {
            const int r = lane & 15;
            float v0 = 0.f, v1 = 0.f, v2 = 0.f;
#pragma unroll
            for (int w8 = 0; w8 < 8; ++w8) {
                const float4 p = *(const float4 *)(rb + (w8 * 16 + r) * 4);
                v0 += p.x;
                v1 += p.y;
                v2 += p.z;
            }
            const float yv = fmaf(16384.0f, v2, fmaf(128.0f, v1, v0));
            const int row = tile * TILE_ROWS + r;
            if (OUT == OUT_BIAS) {
                if (lane < 16 && row < a.rows) a.y[row] = a.bias_or_res ? yv + a.bias_or_res[row] : yv;
            } else if (OUT == OUT_RES) {
                if (lane < 16 && row < a.rows) a.y[row] = __ldcg(a.bias_or_res + row) + yv;
            } else if (OUT == OUT_LOGITS) {
                if (lane < 16 && row < a.rows) {
                    a.y[row] = yv;
                    if (yv > best_v || (yv == best_v && row < best_i)) { best_v = yv; best_i = row; }
                }
            } else {   // OUT_ACTMUL: rows (2j, 2j+1) = (gate_j, up_j)
                const float up = __shfl_down_sync(0xffffffffu, yv, 1);
                const float hv = act_fn(yv, a.act) * up;
                const int pair = tile * 8 + (r >> 1);
                if (lane < 16 && !(r & 1)) {
                    hbuf[(tile & 3) * 8 + (r >> 1)] = hv;
                    if (a.y && 2 * pair < a.rows) a.y[pair] = hv;
                }
                __syncwarp();
                if ((tile & 3) == 3) {   // 32 consecutive h values are complete: put them on the activation grid
                    const int blk = tile >> 2;
                    float sc;
                    const float hx = (2 * (blk * 32 + lane) < a.rows) ? hbuf[lane] : 0.f;
                    int X = act_grid_lane(hx, sc);
                    a3_store_block(a.act_out, a.n_out_act, blk, lane, X, sc);
                }
                __syncwarp();
            }
        }
    }
    if (OUT == OUT_LOGITS && warp == 0) {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, best_v, o);
            int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            if (ov > best_v || (ov == best_v && oi < best_i)) { best_v = ov; best_i = oi; }
        }
        if (lane == 0) { a.pmax[blockIdx.x] = best_v; a.pidx[blockIdx.x] = best_i; }
    }
}

// ------------------------------------------------------------------------------------------ launcher
static int g_num_sms = 0;
int gemv_grid(int n_tiles, int tile_align) {
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    int units = (n_tiles + tile_align - 1) / tile_align;
    return units < g_num_sms ? units : g_num_sms;
}
#define GEMV_FIXED_SMEM (4096 + 192 + 128)
// ring depth: as deep as fits next to the activation image while two CTAs still share an SM (113 KB each)
static int gemv_slots(int nbp_in) {
    const int budget = 113 * 1024 - GEMV_FIXED_SMEM - nbp_in * 100 - 256;
    int s = budget / CHUNK_BYTES;
    if (s > GEMV_MAX_SLOTS) s = GEMV_MAX_SLOTS;
    if (s < 3) s = 3;   // very long inputs: one CTA per SM
    return s;
}

template <int IN, int OUT>
static cudaError_t launch_one(const GemvArgs &a, int grid, size_t smem, cudaStream_t st, bool pdl) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_gemv<IN, OUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
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
    return cudaLaunchKernelEx(&cfg, k_gemv<IN, OUT>, a);
}

// in_mode IN_NORM: (x, norm_w, eps, plus_one) ; IN_ACT: act_in.  out_mode selects the epilogue.
cudaError_t launch_gemv(int in_mode, int out_mode, const void *d_packed, int64_t rows, int64_t cols, const float *x, const float *norm_w,
                        float eps, int plus_one, const void *act_in, float *y, const float *bias_or_res, void *act_out, int act,
                        float *pmax, int *pidx, int *grid_out, cudaStream_t st, bool pdl) {
    GemvArgs a = {};
    a.w = (const uint8_t *)d_packed;
    a.rows = (int)rows;
    a.nbp = (int)nblocks_padded(cols);
    a.n_tiles = (int)ntiles(rows);
    a.tile_align = out_mode == OUT_ACTMUL ? 4 : 1;
    a.n_slots = gemv_slots(a.nbp);
    a.x = x; a.norm_w = norm_w; a.eps = eps; a.plus_one = plus_one; a.n_in = (int)cols;
    a.act_in = (const uint8_t *)act_in;
    a.y = y; a.bias_or_res = bias_or_res; a.act_out = (uint8_t *)act_out; a.n_out_act = (int)(rows / 2); a.act = act;
    a.pmax = pmax; a.pidx = pidx;
    int grid = gemv_grid(a.n_tiles, a.tile_align);
    if (grid_out) *grid_out = grid;
    size_t smem = (size_t)a.n_slots * CHUNK_BYTES + GEMV_FIXED_SMEM + (size_t)a.nbp * 100 + 128;
#define CASE(I, O) if (in_mode == I && out_mode == O) return launch_one<I, O>(a, grid, smem, st, pdl)
    CASE(IN_NORM, OUT_BIAS); CASE(IN_NORM, OUT_RES); CASE(IN_NORM, OUT_ACTMUL); CASE(IN_NORM, OUT_LOGITS);
    CASE(IN_ACT, OUT_BIAS); CASE(IN_ACT, OUT_RES); CASE(IN_ACT, OUT_ACTMUL); CASE(IN_ACT, OUT_LOGITS);
#undef CASE
    return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------------------------------ embedding row
__global__ void k_embed_row(const uint8_t *__restrict__ w, int64_t nbp, int n, const int *__restrict__ token, float scale, float *__restrict__ x) {
    pdl_wait();
    const int64_t row = __ldcg(token);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float v = packed_elem(w, row, i, nbp);
        x[i] = scale != 1.0f ? v * scale : v;
    }
}
void launch_embed_row(const void *d_packed, int64_t cols, const int *d_token, float scale, float *d_x, cudaStream_t st) {
    k_embed_row<<<(unsigned)((cols + 255) / 256), 256, 0, st>>>((const uint8_t *)d_packed, nblocks_padded(cols), (int)cols, d_token, scale, d_x);
}
