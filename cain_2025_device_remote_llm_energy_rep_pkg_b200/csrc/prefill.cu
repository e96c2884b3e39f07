// prefill.cu — batched prompt evaluation on tcgen05 tensor cores (see prefill.cuh for the data layout and the numerics).
//
// k_prefill_gemm: one CTA per SM, persistent over work items (128 weight rows x a K range); fourteen warps in four roles:
//   warps 4-11 stream the packed rows of the group from HBM (the decode layout, layout.cuh: no second copy of the weights),
//              dequantise them exactly to fp32 (Q4_0: d*(q-8); Q4_K: d*sc*q - dmin*m; Q6_K: d*sc*(q-32), each with the oracle's
//              roundings), split each value EXACTLY into NPART bf16 parts (Q4_0: 15 significant bits = 2 parts, K-quants: 24 = 3) and
//              store the parts as A operands
//              (K-major 8 x 16-byte core matrices), one stage = 2 blocks = 64 k = 4 MMA k-steps (128 tokens: 1 block = 2 k-steps)
//   warp 13    copies the B operand of the stage with cp.async.bulk: the three digit planes of the T tokens, already scaled
//              (digit * 2^e * 128^plane is exact in bf16) and already in operand layout
//   warp 12    one thread issues, per k-step and part, ONE tcgen05.mma.kind::f16 (M = 128, N = 3T, K = 16, bf16 x bf16 -> fp32; T = 128:
//              two of N = 192) whose columns are the three digit planes of the T tokens; all parts and k-steps ACCUMULATE into the same
//              columns of tensor memory: every product is exact (8 x 8 significant bits), so the sum over parts and k (and, in the
//              drain, planes) is the fp32 dot product of the dequantised row with the gridded activations
//   warps 0-3  drain the accumulator into fp32 registers once per PF_SEG stages (a bounded chain of tensor-core additions, then one
//              IEEE addition per segment), double-buffered against the MMAs of the next segment; after the last segment they apply
//              the epilogue (+bias, +residual, act(gate)*up) and write y.  T = 128 fills 384 of the 512 columns with ONE accumulator:
//              the whole K range of a work item stays in tensor memory and is drained once, 16 tokens at a time into the epilogue
// Round 1 issued one int8 MMA per 32-element block with accumulation off and combined the digit planes and the two per-block
// scales on the CUDA cores (about ten operations per row, token and block, and a hand-over between the MMA and the epilogue warps
// per block): the tensor pipe was 7-8 % busy.  Folding both scales into the operands keeps the whole K loop in tensor memory.
#include "prefill.cuh"
#include "gemv_dev.cuh"
#include <type_traits>

// ------------------------------------------------------------------------------------------ activation batch -> PB
template <int IN>
__global__ void __launch_bounds__(256) k_prefill_quant(const float *__restrict__ x, int64_t ldx, int M, int T, int n, int nbp,
                                                       const float *__restrict__ norm_w, float eps, int plus_one, uint8_t *__restrict__ pb) {
    __shared__ float red[8];
    __shared__ float s_r;
    const int m = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool live = m < M;
    const float *xr = x + (int64_t)(live ? m : 0) * ldx;
    float r = 1.0f;
    if (IN == PF_IN_NORM && live) {
        float ss = 0.f;
        for (int i = tid; i < n; i += 256) { const float v = xr[i]; ss = fmaf(v, v, ss); }
        ss = warp_sum(ss);
        if (lane == 0) red[warp] = ss;
        __syncthreads();
        if (tid == 0) {
            float tot = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) tot += red[w];
            s_r = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(tot, (float)n), eps)));   // op order of oracle/model.py:_norm
        }
        __syncthreads();
        r = s_r;
    }
    // the blocks of a row are dealt over gridDim.y CTAs (32 CTAs of one token each left most of the GPU idle: 24 us per launch)
    for (int kb = blockIdx.y * 8 + warp; kb < nbp; kb += 8 * gridDim.y) {
        const int i = kb * QK + lane;
        float y = 0.f;
        if (live && i < n) {
            y = xr[i];
            if (IN == PF_IN_NORM) {
                const float w = norm_w[i];
                y = __fmul_rn(__fmul_rn(y, r), plus_one ? __fadd_rn(1.0f, w) : w);
            }
        }
        float sc;
        const int X = act_grid_lane(y, sc);
        int d0, d1, d2;
        act_digits(X, d0, d1, d2);
        // digit * 2^(e-19) * 128^plane: at most 7 significant bits, so the fp32 product is exact and its upper half is the bf16 value
        const float v0 = (float)d0 * sc, v1 = (float)d1 * (sc * 128.0f), v2 = (float)d2 * (sc * 16384.0f);
        *(uint16_t *)(pb + pb_off(T, kb, 0, m, lane)) = (uint16_t)(__float_as_uint(v0) >> 16);
        *(uint16_t *)(pb + pb_off(T, kb, 1, m, lane)) = (uint16_t)(__float_as_uint(v1) >> 16);
        *(uint16_t *)(pb + pb_off(T, kb, 2, m, lane)) = (uint16_t)(__float_as_uint(v2) >> 16);
    }
}

cudaError_t launch_prefill_quant(int in_mode, const float *x, int64_t ldx, int M, int T, int64_t n, const float *norm_w, float eps, int plus_one,
                                 void *pb, cudaStream_t st) {
    if (M <= 0 || M > T || (T != 16 && T != 32 && T != 64 && T != 128) || n <= 0) return cudaErrorInvalidValue;
    const int nbp = (int)nblocks_padded(n);
    const int slices = nbp >= 256 ? 8 : nbp >= 64 ? 4 : 1;   // every CTA of a token recomputes the row's sum of squares (a few KB from L2)
    if (in_mode == PF_IN_NORM) k_prefill_quant<PF_IN_NORM><<<dim3(T, slices), 256, 0, st>>>(x, ldx, M, T, (int)n, nbp, norm_w, eps, plus_one, (uint8_t *)pb);
    else k_prefill_quant<PF_IN_F32><<<dim3(T, slices), 256, 0, st>>>(x, ldx, M, T, (int)n, nbp, nullptr, 0.f, 0, (uint8_t *)pb);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ tcgen05 helpers
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst, uint32_t cols) {   // one whole warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {      // the same warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], bf16 x bf16 -> fp32, M = 128, K = 16; accumulate = 0 overwrites D
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// the mbarrier gets one arrival when every MMA this thread has issued so far has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 16 consecutive 32-bit columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld16(uint32_t addr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(addr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// shared-memory matrix descriptor: K-major, no swizzle; core matrix = 8 rows x 16 bytes (128 contiguous bytes), the two 16-byte
// K halves of a 32-byte row (16 bf16) 128 bytes apart (leading-dimension offset), 8-row groups 256 bytes apart (stride offset)
__device__ __forceinline__ uint64_t umma_desc(const void *smem) {
    return (uint64_t)((smem_u32(smem) >> 4) & 0x3FFF) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46);
}
// instruction descriptor of kind::f16: D fp32, A and B bf16, both K-major, N = n, M = 128
__host__ __device__ inline uint32_t umma_idesc_bf16(int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(PF_ROWS >> 4) << 24);
}

struct PfArgs {
    const uint8_t *w;        // packed Q4_0 / Q4_K matrix (layout.cuh)
    int rows, nbp, n_tiles;
    const uint8_t *pb;       // B operand of the batch (prefill.cuh)
    int M;                   // live tokens (<= T)
    float *y;
    int64_t ldy;
    const float *bias_or_res;
    int act;
    int force_part;          // 1: raw sums go to `part` even when ksplit is 1 (tensor parallel: the sums are exchanged, not finished here)
    int ksplit;              // > 1: the K range of a row group is cut into ksplit pieces, one work item each; raw sums go to `part`
    float *part;             // [ksplit][T][rows] fp32 partial sums (k_prefill_finish adds them in a fixed order)
};

#ifndef PF_STAGE_BLOCKS
#define PF_STAGE_BLOCKS 4   // 32-element blocks per stage: 4 (a whole block-group) where two such stages fit, else 2; 1 and 2 force smaller stages
#endif
// Measured, llama3.1:8b, ms per prompt with stages of 1 / 2 / 4 blocks: 16 tokens 8.8 / 8.1 / 7.7, 32 tokens 9.2 / 8.7 / 8.2, 64 tokens 10.4 / 9.7 / 9.3,
// 128 tokens 14.5 / 13.2 / - (five stages of 40 KB against two of 80 KB), 512 tokens 62.1 / 57.0 / -.  What a stage costs beyond its arithmetic
// (barrier round trips, the proxy fence, the commit, the MMA thread's wake-up) is per stage, and a deeper ring buys nothing back; one fence
// per two one-block stages changed nothing (14.5).
// shared memory: [A stages: STAGES x (2 SB k-steps x NPART images of 128 rows x 32 bytes)][B stages: STAGES x (2 SB k-steps x 3T rows x 32 bytes)][barriers]
template <int T, int NPART>
struct PfSmem {
    static constexpr int stage_bytes(int sb) { return 2 * sb * NPART * PF_ROWS * 32 + 2 * sb * 3 * T * 32; }
    // blocks (32 k = 2 k-steps) per stage; a whole block-group (4) only where two such stages fit
    static constexpr int SB = (PF_STAGE_BLOCKS == 4 && (227 * 1024 - 256) / stage_bytes(4) >= 2) ? 4 : PF_STAGE_BLOCKS >= 2 ? 2 : 1;
    // as many stages as fit: the round trip of a stage (MMAs done -> rebuilt -> MMAs issued again) is several times the work of a stage
    static constexpr int FIT = (227 * 1024 - 256) / stage_bytes(SB);
    static constexpr int STAGES = FIT > 6 ? 6 : FIT;
    static constexpr int A_IMG = PF_ROWS * 32;                 // one A operand: 128 rows x 16 bf16
    static constexpr int A_STAGE = 2 * SB * NPART * A_IMG;
    static constexpr int B_KSTEP = 3 * T * 32;                 // one k-step of B: 3 planes x T tokens x 16 bf16
    static constexpr int B_STAGE = 2 * SB * B_KSTEP;           // = SB blocks of the PB image
    static constexpr int A_OFF = 0;
    static constexpr int B_OFF = A_OFF + STAGES * A_STAGE;
    static constexpr int BAR_OFF = B_OFF + STAGES * B_STAGE;
    static constexpr int TOTAL = BAR_OFF + 256;
};

// one fp32 value as NPART bf16 parts whose sum is the value (truncation leaves a remainder with 8 significant bits less)
template <int NPART>
__device__ __forceinline__ void split_bf16(float w, uint32_t (&part)[NPART]) {
#pragma unroll
    for (int p = 0; p < NPART; ++p) {
        part[p] = __float_as_uint(w) & 0xFFFF0000u;
        if (p + 1 < NPART) w = __fsub_rn(w, __uint_as_float(part[p]));
    }
}

#ifndef PF_RING3
#define PF_RING3 0          // 1: three register sets of weights rotate through a loop unrolled by three (two block-groups ahead)
#endif
#ifndef PF_HS_UNROLL
#define PF_HS_UNROLL 2      // 1: the two stages of a block-group share one copy of the dequantisation code
#endif
template <int T, int OUT, int QT>
__global__ void __launch_bounds__(PF_THREADS, 1) k_prefill_gemm(const PfArgs a) {
    constexpr int NPART = QT ? 3 : 2;
    // One MMA per (k-step, part) with N = 3T: the three digit planes are columns [plane*T + token] of the same MMA, so an A image
    // (4 KB) is fetched from shared memory once per k-step and part instead of once per plane.  Measured with N = T and three MMAs
    // per part: 14 ms per pass over llama3.1:8b whatever T was, because every small MMA re-reads its whole A operand and the
    // operand fetch (24 x 4 KB per 64 k) was what the tensor pipe waited for.  The parts alternate between NACC accumulators.
    constexpr int N = 3 * T;
    constexpr int NACC = T <= 32 ? 2 : 1;              // 2 buffers x NACC x 3T columns <= 384 of the 512 columns of tensor memory
    // 128 tokens: N = 384 is two MMAs of 192 columns per part and k-step, and the 384 columns exist once.  The sums of a work item's whole
    // K range then stay in tensor memory (one segment) and are drained once, 16 tokens at a time straight into the epilogue: a thread
    // cannot hold 128 running sums.  Measured on same-sign data, K = 14 336: 1.3e-6 of sum |w||x| against 2e-7 with 512-k segments.
    constexpr int NM = N > 256 ? 2 : 1, NMMA = N / NM;
    constexpr int NBUF = 2 * NACC * N <= 512 ? 2 : 1;
    constexpr bool ONE_SEG = NBUF == 1;
    using S = PfSmem<T, NPART>;
    constexpr int PF_STAGES = S::STAGES, SB = S::SB, ST_PER_BG = 4 / SB;   // a block-group (4 blocks) is ST_PER_BG stages
    constexpr int SEG = ONE_SEG ? (1 << 30) : PF_SEG * 2 / SB;   // PF_SEG counts two-block stages
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t *full = (uint64_t *)(smem + S::BAR_OFF);   // [stage] A built (8 warp arrivals) + B copied (1 arrival + bytes)
    uint64_t *empty = full + PF_STAGES;                 // [stage] the MMAs that read the stage are done (commit)
    uint64_t *tfull = empty + PF_STAGES;                // [buf]   a segment's sums are complete in tensor memory
    uint64_t *tempty = tfull + 2;                       // [buf]   the four epilogue warps have drained them
    uint32_t *tmem_ptr = (uint32_t *)(tempty + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < PF_STAGES; ++s) { mbar_init(full + s, 9); mbar_init(empty + s, 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(tfull + b, 1); mbar_init(tempty + b, 4); }   // NBUF of them are used
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 12) tmem_alloc(tmem_ptr, 512);          // NBUF buffers of NACC accumulators of 3T columns
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_ptr;
    const int n_items = ((a.n_tiles + 7) >> 3) * a.ksplit;   // work item = (row group, K piece)
    const int nbg_all = a.nbp >> 2;                          // block-groups along K; a block-group is ST_PER_BG stages
    auto item_range = [&](int wi, int &grp, int &ks, int &bg0, int &bg1) {
        grp = wi / a.ksplit;
        ks = wi - grp * a.ksplit;
        bg0 = (int)(((int64_t)nbg_all * ks) / a.ksplit);
        bg1 = (int)(((int64_t)nbg_all * (ks + 1)) / a.ksplit);
    };

    if (warp < 4) {
        // ===================== accumulator drain + epilogue: thread = weight row (TMEM lane), T fp32 sums
        const int row_l = warp * 32 + lane;
        const uint32_t tlane = tmem + ((uint32_t)(warp * 32) << 16);
        int seg = 0;
        for (int wi = blockIdx.x; wi < n_items; wi += gridDim.x) {
            int grp, ks, bg0, bg1;
            item_range(wi, grp, ks, bg0, bg1);
            const int n_seg = ONE_SEG ? 1 : (ST_PER_BG * (bg1 - bg0) + SEG - 1) / SEG;
            const int row = grp * PF_ROWS + row_l;
            // the finished sums of tokens t0 .. t0+15 of this thread's row
            auto emit16 = [&](int t0, const float (&v)[16]) {
                if (a.ksplit > 1 || a.force_part) {
#pragma unroll
                    for (int u = 0; u < 16; ++u)
                        if (t0 + u < a.M && row < a.rows) a.part[((int64_t)ks * T + t0 + u) * a.rows + row] = v[u];
                    return;
                }
#pragma unroll
                for (int u = 0; u < 16; ++u) {
                    const int tk = t0 + u;
                    if (tk < a.M) {   // uniform
                        if (OUT == PF_OUT_ACTMUL) {   // rows (2j, 2j+1) = (gate_j, up_j)
                            const float up = __shfl_down_sync(0xffffffffu, v[u], 1);
                            if (!(row & 1) && row < a.rows) a.y[(int64_t)tk * a.ldy + (row >> 1)] = act_fn(v[u], a.act) * up;
                        } else if (row < a.rows) {
                            if (OUT == PF_OUT_RES) a.y[(int64_t)tk * a.ldy + row] = a.bias_or_res[(int64_t)tk * a.ldy + row] + v[u];
                            else a.y[(int64_t)tk * a.ldy + row] = a.bias_or_res ? v[u] + a.bias_or_res[row] : v[u];
                        }
                    }
                }
            };
            if constexpr (ONE_SEG) {
                mbar_wait(tfull, seg & 1);
                tc_fence_after();
#pragma unroll 1
                for (int t0 = 0; t0 < T; t0 += 16) {
                    uint32_t c0[16], c1[16], c2[16];
                    tmem_ld16(tlane + t0, c0);
                    tmem_ld16(tlane + T + t0, c1);
                    tmem_ld16(tlane + 2 * T + t0, c2);
                    tmem_ld_wait();
                    if (t0 + 16 >= T) {   // everything is in registers: the MMAs of the next item may overwrite the columns
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(tempty);
                    }
                    float v[16];
#pragma unroll
                    for (int u = 0; u < 16; ++u) v[u] = __fadd_rn(__fadd_rn(__uint_as_float(c0[u]), __uint_as_float(c1[u])), __uint_as_float(c2[u]));
                    emit16(t0, v);
                }
                ++seg;
            } else {
                float acc[T];
#pragma unroll
                for (int m = 0; m < T; ++m) acc[m] = 0.f;
                for (int sg = 0; sg < n_seg; ++sg, ++seg) {
                    const int buf = seg & 1;
                    mbar_wait(tfull + buf, (seg >> 1) & 1);
                    tc_fence_after();
#pragma unroll
                    for (int ai = 0; ai < NACC; ++ai) {
#pragma unroll
                        for (int t0 = 0; t0 < T; t0 += 16) {
                            uint32_t c0[16], c1[16], c2[16];   // the three digit planes of 16 tokens (already scaled by 128^plane)
                            tmem_ld16(tlane + (buf * NACC + ai) * N + t0, c0);
                            tmem_ld16(tlane + (buf * NACC + ai) * N + T + t0, c1);
                            tmem_ld16(tlane + (buf * NACC + ai) * N + 2 * T + t0, c2);
                            tmem_ld_wait();
#pragma unroll
                            for (int u = 0; u < 16; ++u)
                                acc[t0 + u] = __fadd_rn(acc[t0 + u], __fadd_rn(__fadd_rn(__uint_as_float(c0[u]), __uint_as_float(c1[u])), __uint_as_float(c2[u])));
                        }
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(tempty + buf);
                }
#pragma unroll
                for (int t0 = 0; t0 < T; t0 += 16) {
                    float v[16];
#pragma unroll
                    for (int u = 0; u < 16; ++u) v[u] = acc[t0 + u];
                    emit16(t0, v);
                }
            }
        }
    } else if (warp < 12) {
        // ===================== weights: packed nibbles -> exact fp32 -> NPART bf16 A operands.  Warp lw owns tile lw of the group
        // (16 rows); lane (g, t) holds word t (elements 4t..4t+3 in the low nibbles, 16+4t.. in the high ones) of four consecutive
        // blocks of tile rows g and g+8 (two LDG.128 per block-group, as the decode GEMV reads them).
        const int lw = warp - 4, g = lane >> 2, t = lane & 3;
        constexpr int BB = QT == 2 ? BLOCK_BYTES_6 : QT ? BLOCK_BYTES_K : BLOCK_BYTES, SCB = QT ? 192 : 128;
        // the weights of a work item go HBM -> L2 one item ahead (lane 0, one bulk prefetch per chunk), so the loads below are L2 hits
        auto prefetch_item = [&](int wi) {
            if (lane != 0 || wi >= n_items) return;
            int grp, ks, bg0, bg1;
            item_range(wi, grp, ks, bg0, bg1);
            const int tg = grp * 8 + lw;
            if (tg >= a.n_tiles || bg1 <= bg0) return;
            for (int ck = bg0 >> 4; ck <= (bg1 - 1) >> 4; ++ck) {
                const int cb = min(KC, a.nbp - ck * KC);
                l2_prefetch(a.w + ((uint64_t)tg * a.nbp + (uint64_t)ck * KC) * BB, (uint32_t)cb * BB);
            }
        };
        // one block-group of the tile in registers: nibble words of rows g / g+8, this lane's scale word, (d, dmin) of both rows
        // Q6_K adds the high bits of the same elements (wh), keeps the scale words of all four blocks of its rows in the lane (sd4: int8
        // (sc_lo_g, sc_hi_g, sc_lo_g+8, sc_hi_g+8) per block) and half2 (d_g, d_g+8) in sc
        struct Regs4 { uint4 v0, v1; uint32_t sc; uint2 dm; };
        struct Regs6 { uint4 v0, v1, wh, sd4; uint32_t sc; };
        using Regs = typename std::conditional<QT == 2, Regs6, Regs4>::type;
        auto issue = [&](int grp, int bgi, Regs &r) {
            const int ck = bgi >> 4, j = bgi & 15, tg = grp * 8 + lw;
            const int cb = min(KC, a.nbp - ck * KC);
            r.v0 = r.v1 = make_uint4(0u, 0u, 0u, 0u);
            r.sc = 0u;
            if constexpr (QT == 2) { r.wh = r.sd4 = make_uint4(0u, 0u, 0u, 0u); } else { r.dm = make_uint2(0u, 0u); }
            if (tg < a.n_tiles) {
                const uint8_t *chunk = a.w + ((uint64_t)tg * a.nbp + (uint64_t)ck * KC) * BB;
                r.v0 = ldg_stream4(chunk + (size_t)j * 1024 + lane * 16);
                r.v1 = ldg_stream4(chunk + (size_t)j * 1024 + 512 + lane * 16);
                if constexpr (QT == 2) {
                    r.wh = ldg_stream4(chunk + (size_t)cb * 256 + (size_t)j * 512 + lane * 16);
                    r.sd4 = ldg_stream4(chunk + (size_t)cb * 384 + j * 160 + (lane >> 2) * 16);
                    r.sc = __ldg((const uint32_t *)(chunk + (size_t)cb * 384 + j * 160 + 128 + (lane >> 2) * 4));
                } else {
                    // Q4_0: half2 (d of row g, d of row g+8) of block t; Q4_K: bytes (sc_g, m_g, sc_g+8, m_g+8) of block t and half2 (d, dmin) of both rows
                    r.sc = __ldg((const uint32_t *)(chunk + (size_t)cb * 256 + j * SCB + lane * 4));
                    if constexpr (QT == 1) r.dm = __ldg((const uint2 *)(chunk + (size_t)cb * 256 + j * SCB + 128 + (lane >> 2) * 8));
                }
            }
        };
        // The loads of the next block-group are in flight while the current one is turned into A operands.  How far ahead does not
        // matter (0 / 1 block-groups ahead: 9.4 / 9.3 ms per 32-token prompt; the loads hit L2 behind the item-ahead bulk prefetch), how
        // much CODE the loop has does: ncu shows 21-33 % of the warp samples of this kernel waiting for instructions.  Two block-groups
        // ahead through a loop unrolled by three (three copies of the dequantisation code, -DPF_RING3=1) costs 14 %: 9.8 / 11.1 / 96 ms
        // against 8.6 / 9.7 / 85 ms for prompts of 32 / 64 / 512 tokens with the single copy below.
        struct Cursor { int wi, grp, ks, bgi, bg1; };
        auto start = [&](Cursor &c, int wi) {
            c.wi = wi; c.grp = c.ks = c.bgi = c.bg1 = 0;
            if (wi < n_items) item_range(wi, c.grp, c.ks, c.bgi, c.bg1);
        };
        auto advance = [&](Cursor &c, bool prefetch) {   // to the next block-group, possibly the first of the CTA's next work item
            if (c.wi >= n_items) return;
            if (++c.bgi >= c.bg1) {
                start(c, c.wi + gridDim.x);
                if (prefetch) prefetch_item(c.wi + gridDim.x);
            }
        };
        int it = 0;
        Cursor use, fill;
        start(use, blockIdx.x);
        fill = use;
        prefetch_item(use.wi);
        prefetch_item(use.wi + gridDim.x);
        Regs r0, r1, r2;
        if (fill.wi < n_items) issue(fill.grp, fill.bgi, r0);
        advance(fill, true);
#if PF_RING3
        if (fill.wi < n_items) issue(fill.grp, fill.bgi, r1);
        advance(fill, true);
#endif
        auto build = [&](const Regs &cur) {
            const uint32_t w0[4] = {cur.v0.x ^ 0x88888888u, cur.v0.y ^ 0x88888888u, cur.v0.z ^ 0x88888888u, cur.v0.w ^ 0x88888888u};   // GGML nibbles q of row g
            const uint32_t w1[4] = {cur.v1.x ^ 0x88888888u, cur.v1.y ^ 0x88888888u, cur.v1.z ^ 0x88888888u, cur.v1.w ^ 0x88888888u};   // row g+8
            constexpr int HS_UNROLL = SB == 2 ? PF_HS_UNROLL : SB == 1 ? 4 : 1;   // the block index jj must be a constant: it selects registers
#pragma unroll HS_UNROLL
            for (int hs = 0; hs < ST_PER_BG; ++hs, ++it) {   // the stages of the block-group: blocks SB*hs .. SB*hs + SB-1
                const int slot = it % PF_STAGES;
                mbar_wait(empty + slot, ((it / PF_STAGES) & 1) ^ 1);
                uint8_t *As = smem + S::A_OFF + slot * S::A_STAGE;
#pragma unroll
                for (int bb = 0; bb < SB; ++bb) {
                    const int jj = SB * hs + bb;
                    // mul[row half][k-step], sub[row half]: Q4_0 w = mul*(q-8); Q4_K w = q*mul - sub; Q6_K w = mul*(q-32), mul per 16 elements
                    float mul[2][2], sub[2];
                    uint32_t hbits[2] = {0u, 0u};   // Q6_K: the 16-bit field of high bits of (row half, block jj): byte kh, 2 bits per element
                    if constexpr (QT == 0) {
                        const uint32_t scw = __shfl_sync(0xffffffffu, cur.sc, g * 4 + jj);   // the scale word of (rows g / g+8, block jj)
                        const float2 d2 = __half22float2(*(const __half2 *)&scw);
                        mul[0][0] = mul[0][1] = d2.x; mul[1][0] = mul[1][1] = d2.y;
                        sub[0] = sub[1] = 0.f;
                    } else if constexpr (QT == 1) {   // d1 = d*sc, m1 = dmin*m, each rounded (gguf-py / oracle/quants.py order)
                        const uint32_t scw = __shfl_sync(0xffffffffu, cur.sc, g * 4 + jj);
                        const float2 dg = __half22float2(*(const __half2 *)&cur.dm.x), dg8 = __half22float2(*(const __half2 *)&cur.dm.y);
                        mul[0][0] = mul[0][1] = __fmul_rn(dg.x, byte_to_float(scw, 0)); sub[0] = __fmul_rn(dg.y, byte_to_float(scw, 1));
                        mul[1][0] = mul[1][1] = __fmul_rn(dg8.x, byte_to_float(scw, 2)); sub[1] = __fmul_rn(dg8.y, byte_to_float(scw, 3));
                    } else {   // dsc = d*sc per 16-element half, rounded (oracle/quants.py order); int8 scale as byte + 128
                        const uint32_t sw = (jj == 0 ? cur.sd4.x : jj == 1 ? cur.sd4.y : jj == 2 ? cur.sd4.z : cur.sd4.w) ^ 0x80808080u;
                        const float2 dd = __half22float2(*(const __half2 *)&cur.sc);
                        mul[0][0] = __fmul_rn(dd.x, byte_to_float(sw, 0) - 128.0f); mul[0][1] = __fmul_rn(dd.x, byte_to_float(sw, 1) - 128.0f);
                        mul[1][0] = __fmul_rn(dd.y, byte_to_float(sw, 2) - 128.0f); mul[1][1] = __fmul_rn(dd.y, byte_to_float(sw, 3) - 128.0f);
                        sub[0] = sub[1] = 0.f;
                        const uint32_t hg = jj < 2 ? cur.wh.x : cur.wh.y, hg8 = jj < 2 ? cur.wh.z : cur.wh.w;
                        hbits[0] = (hg >> (16 * (jj & 1))) & 0xFFFFu;
                        hbits[1] = (hg8 >> (16 * (jj & 1))) & 0xFFFFu;
                    }
#pragma unroll
                    for (int hf = 0; hf < 2; ++hf) {          // tile row g (hf 0) / g+8 (hf 1)
                        const uint32_t word = hf ? w1[jj] : w0[jj];
#pragma unroll
                        for (int kh = 0; kh < 2; ++kh) {      // k-step 0: elements 4t..4t+3 (low nibbles), k-step 1: 16+4t.. (high nibbles)
                            uint32_t parts[4][NPART];
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                uint32_t q = (word >> (8 * e + 4 * kh)) & 0xFu;
                                if constexpr (QT == 2) q |= ((hbits[hf] >> (8 * kh + 2 * e)) & 3u) << 4;
                                // q as a float without I2F: 0x4B000000 | q is 8388608 + q; the offset of the format goes into the same subtraction
                                const float qf = __uint_as_float(0x4B000000u | q) - (QT == 0 ? 8388616.0f : QT == 1 ? 8388608.0f : 8388640.0f);
                                const float wv = QT == 1 ? __fsub_rn(__fmul_rn(qf, mul[hf][kh]), sub[hf]) : __fmul_rn(mul[hf][kh], qf);
                                split_bf16<NPART>(wv, parts[e]);
                            }
                            // 4 consecutive k of row (hf*8 + g) of tile lw: 8 bytes of the image of k-step (bb*2 + kh)
                            uint8_t *dst = As + ((bb * 2 + kh) * NPART) * S::A_IMG + (lw * 2 + hf) * 256 + (t >> 1) * 128 + g * 16 + (t & 1) * 8;
#pragma unroll
                            for (int p = 0; p < NPART; ++p)
                                *(uint2 *)(dst + p * S::A_IMG) = make_uint2(__byte_perm(parts[0][p], parts[1][p], 0x7632), __byte_perm(parts[2][p], parts[3][p], 0x7632));
                        }
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the tensor core reads these stores through the async proxy
                __syncwarp();
                if (lane == 0) mbar_arrive(full + slot);
            }
        };
        auto step = [&](const Regs &cur, Regs &spare) {   // spare <- the block-group two ahead; cur -> shared memory
            if (use.wi >= n_items) return;
            if (fill.wi < n_items) issue(fill.grp, fill.bgi, spare);
            advance(fill, true);
            build(cur);
            advance(use, false);
        };
#if PF_RING3
        while (use.wi < n_items) {
            step(r0, r2);
            step(r1, r0);
            step(r2, r1);
        }
#else
        (void)r2;
        while (use.wi < n_items) {   // one block-group ahead, one copy of the build code
            if (fill.wi < n_items) issue(fill.grp, fill.bgi, r1);
            advance(fill, true);
            build(r0);
            advance(use, false);
            r0 = r1;
        }
#endif
    } else if (warp == 12) {
        // ===================== MMA issue (one thread)
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(NMMA);
            // descriptors by addition: the start-address field (bits 0-13, in 16-byte units) of a descriptor of the first image plus
            // the image's offset; one thread issues everything here, so every instruction it saves shortens the stage
            const uint64_t descA0 = umma_desc(smem + S::A_OFF), descB0 = umma_desc(smem + S::B_OFF);
            int it = 0, seg = 0;
            for (int wi = blockIdx.x; wi < n_items; wi += gridDim.x) {
                int grp, ks, bg0, bg1;
                item_range(wi, grp, ks, bg0, bg1);
                const int n_st = ST_PER_BG * (bg1 - bg0);
                for (int st = 0; st < n_st; ++st, ++it) {
                    const int slot = it % PF_STAGES, in_seg = ONE_SEG ? st : st % SEG, buf = seg % NBUF;
                    if (in_seg == 0) {   // a new segment: its accumulator buffer must have been drained
                        mbar_wait(tempty + buf, ((seg / NBUF) & 1) ^ 1);
                        tc_fence_after();
                    }
                    mbar_wait(full + slot, (it / PF_STAGES) & 1);
                    tc_fence_after();
                    const uint64_t da_s = descA0 + (uint64_t)((slot * S::A_STAGE) >> 4), db_s = descB0 + (uint64_t)((slot * S::B_STAGE) >> 4);
                    const uint32_t acc_on = in_seg ? 1u : 0u;
#pragma unroll
                    for (int kst = 0; kst < 2 * SB; ++kst) {
                        const uint64_t db = db_s + (uint64_t)((kst * S::B_KSTEP) >> 4);
#pragma unroll
                        for (int p = 0; p < NPART; ++p) {
                            const int ai = p % NACC;     // the first MMA into an accumulator of a segment overwrites it
#pragma unroll
                            for (int j = 0; j < NM; ++j)   // N > 256: columns [j*NMMA, (j+1)*NMMA) = rows of the B image 256 bytes per 8 apart
                                umma_bf16(tmem + (buf * NACC + ai) * N + j * NMMA, da_s + (uint64_t)(((kst * NPART + p) * S::A_IMG) >> 4),
                                          db + (uint64_t)((j * (NMMA / 8) * 256) >> 4), idesc, (kst || p >= NACC) ? 1u : acc_on);
                        }
                    }
                    umma_commit(empty + slot);
                    if ((!ONE_SEG && in_seg == SEG - 1) || st == n_st - 1) { umma_commit(tfull + buf); ++seg; }
                }
            }
        }
    } else {
        // ===================== B operand of the stage: one bulk copy (SB blocks of the PB image)
        if (lane == 0) {
            int it = 0;
            for (int wi = blockIdx.x; wi < n_items; wi += gridDim.x) {
                int grp, ks, bg0, bg1;
                item_range(wi, grp, ks, bg0, bg1);
                for (int st = ST_PER_BG * bg0; st < ST_PER_BG * bg1; ++st, ++it) {
                    const int slot = it % PF_STAGES;
                    mbar_wait(empty + slot, ((it / PF_STAGES) & 1) ^ 1);
                    mbar_expect_tx(full + slot, (uint32_t)S::B_STAGE);
                    bulk_g2s(smem + S::B_OFF + slot * S::B_STAGE, a.pb + (size_t)st * S::B_STAGE, (uint32_t)S::B_STAGE, full + slot);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 12) tmem_dealloc(tmem, 512);
}

// y = sum of the K-split partials in a fixed order, then the epilogue (+bias | +residual)
__global__ void __launch_bounds__(256) k_prefill_finish(const float *__restrict__ part, int ksplit, int T, int M, int rows, float *__restrict__ y,
                                                        int64_t ldy, const float *__restrict__ bias_or_res, int out_mode) {
    const int row = blockIdx.x * 256 + threadIdx.x, m = blockIdx.y;
    if (row >= rows || m >= M) return;
    float v = 0.f;
    for (int s = 0; s < ksplit; ++s) v += part[((int64_t)s * T + m) * rows + row];
    if (out_mode == PF_OUT_RES) v += bias_or_res[(int64_t)m * ldy + row];
    else if (bias_or_res) v += bias_or_res[row];
    y[(int64_t)m * ldy + row] = v;
}
// the same for the gate/up GEMM: rows (2j, 2j+1) = (gate_j, up_j), y[m][j] = act(gate) * up
__global__ void __launch_bounds__(256) k_prefill_finish_actmul(const float *__restrict__ part, int ksplit, int T, int M, int rows, float *__restrict__ y,
                                                               int64_t ldy, int act) {
    const int j = blockIdx.x * 256 + threadIdx.x, m = blockIdx.y;
    if (2 * j + 1 >= rows || m >= M) return;
    float g = 0.f, u = 0.f;
    for (int s = 0; s < ksplit; ++s) {
        const float2 gu = *(const float2 *)(part + ((int64_t)s * T + m) * rows + 2 * j);
        g += gu.x;
        u += gu.y;
    }
    y[(int64_t)m * ldy + j] = act_fn(g, act) * u;
}

template <int T, int OUT, int QT>
static cudaError_t launch_pf_qt(const PfArgs &a, int grid, cudaStream_t st) {
    using S = PfSmem<T, QT ? 3 : 2>;
    static_assert(S::TOTAL <= 227 * 1024, "prompt GEMM: shared memory");
    static bool attr_set[B2_MAX_DEVICES] = {};   // function attributes belong to a device (one engine per GPU in one server process)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= B2_MAX_DEVICES || !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(k_prefill_gemm<T, OUT, QT>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < B2_MAX_DEVICES) attr_set[dev] = true;
    }
    k_prefill_gemm<T, OUT, QT><<<grid, PF_THREADS, S::TOTAL, st>>>(a);
    return cudaGetLastError();
}
template <int T, int OUT>
static cudaError_t launch_pf(const PfArgs &a, int qt, int grid, cudaStream_t st) {
    return qt == 2 ? launch_pf_qt<T, OUT, 2>(a, grid, st) : qt ? launch_pf_qt<T, OUT, 1>(a, grid, st) : launch_pf_qt<T, OUT, 0>(a, grid, st);
}

// number of K pieces for n_groups row groups on sm_count CTAs: the smallest count whose launch time, ceil(groups * pieces / CTAs) / pieces
// item times, is within 5 % of the best reachable with at most kmax pieces
int pf_pick_ksplit(int n_groups, int sm_count, int kmax) {
    int ksplit = 1;
    if (n_groups <= 0 || sm_count <= 0) return 1;
    double best = (double)((n_groups + sm_count - 1) / sm_count);
    for (int ks = 2; ks <= kmax; ++ks) {
        const double t = (double)(((int64_t)n_groups * ks + sm_count - 1) / sm_count) / ks;
        if (t < 0.95 * best) { best = t; ksplit = ks; }
    }
    return ksplit;
}
extern "C" int b2dbg_pf_ksplit(int n_groups, int sm_count, int kmax) { return pf_pick_ksplit(n_groups, sm_count, kmax); }   // host logic, tested on CPU

cudaError_t launch_prefill_gemm(int out_mode, const void *d_packed, int64_t rows, int64_t cols, const void *pb, int M, int T, float *y, int64_t ldy,
                                const float *bias_or_res, int act, int sm_count, float *part, uint64_t part_floats, cudaStream_t st) {
    return launch_prefill_gemm_qt(0, out_mode, d_packed, rows, cols, pb, M, T, y, ldy, bias_or_res, act, sm_count, part, part_floats, st);
}
cudaError_t launch_prefill_gemm_qt(int qt, int out_mode, const void *d_packed, int64_t rows, int64_t cols, const void *pb, int M, int T, float *y, int64_t ldy,
                                   const float *bias_or_res, int act, int sm_count, float *part, uint64_t part_floats, cudaStream_t st,
                                   bool partials_only, int *ksplit_out) {
    if (M <= 0 || M > T || rows <= 0 || cols <= 0) return cudaErrorInvalidValue;
    if (partials_only && (!part || out_mode == PF_OUT_ACTMUL || (uint64_t)T * rows > part_floats)) return cudaErrorInvalidValue;
    if (out_mode == PF_OUT_RES && !bias_or_res) return cudaErrorInvalidValue;
    PfArgs a;
    a.w = (const uint8_t *)d_packed; a.rows = (int)rows; a.nbp = (int)nblocks_padded(cols); a.n_tiles = (int)ntiles(rows);
    a.pb = (const uint8_t *)pb; a.M = M; a.y = y; a.ldy = ldy; a.bias_or_res = bias_or_res; a.act = act;
    const int n_groups = (a.n_tiles + 7) / 8;
    // Work items are dealt round robin to one CTA per SM, so a launch takes ceil(items / SMs) item times.  Matrices with few row groups
    // (O, down, QKV) would leave most SMs idle, and 224 row groups (llama3.1:8b gate/up) on 148 SMs take two rounds for 1.5 rounds of
    // work: K is cut into the smallest number of pieces that brings the launch within 5 % of the best reachable (224 groups: 3 pieces,
    // 672 items = 5 rounds of a third).  The pieces' raw sums go to `part` and a second kernel adds them in a fixed order.
    int ksplit = 1;
    if (part && (out_mode != PF_OUT_ACTMUL || (rows & 1) == 0)) {
        const int nbg = a.nbp >> 2;
        int kmax = nbg / 8 < 8 ? nbg / 8 : 8;   // at least 8 block-groups per item
        const uint64_t per_piece = (uint64_t)T * rows;
        if (per_piece * (uint64_t)kmax > part_floats) kmax = (int)(part_floats / per_piece);
        ksplit = pf_pick_ksplit(n_groups, sm_count, kmax);
    }
    a.ksplit = ksplit; a.part = part; a.force_part = partials_only ? 1 : 0;
    if (ksplit_out) *ksplit_out = ksplit;
    const int n_items = n_groups * ksplit;
    const int grid = n_items < sm_count ? n_items : sm_count;
    cudaError_t e = cudaErrorInvalidValue;
#define PF_CASE(TT, OO) if (T == TT && out_mode == OO) e = launch_pf<TT, OO>(a, qt, grid, st)
    PF_CASE(16, PF_OUT_BIAS); PF_CASE(16, PF_OUT_RES); PF_CASE(16, PF_OUT_ACTMUL);
    PF_CASE(32, PF_OUT_BIAS); PF_CASE(32, PF_OUT_RES); PF_CASE(32, PF_OUT_ACTMUL);
    PF_CASE(64, PF_OUT_BIAS); PF_CASE(64, PF_OUT_RES); PF_CASE(64, PF_OUT_ACTMUL);
    PF_CASE(128, PF_OUT_BIAS); PF_CASE(128, PF_OUT_RES); PF_CASE(128, PF_OUT_ACTMUL);
#undef PF_CASE
    if (e != cudaSuccess || ksplit == 1 || partials_only) return e;
    if (out_mode == PF_OUT_ACTMUL) k_prefill_finish_actmul<<<dim3((unsigned)((rows / 2 + 255) / 256), M), 256, 0, st>>>(part, ksplit, T, M, (int)rows, y, ldy, act);
    else k_prefill_finish<<<dim3((unsigned)((rows + 255) / 256), M), 256, 0, st>>>(part, ksplit, T, M, (int)rows, y, ldy, bias_or_res, out_mode);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ the other stages of a prompt batch
// token embedding rows of M tokens (gemma: * sqrt(H)); grid (M)
template <int QT>
__global__ void __launch_bounds__(256) k_prefill_embed(const uint8_t *__restrict__ w, int64_t nbp, int n, const int *__restrict__ tokens, float scale,
                                                       float *__restrict__ x) {
    const int m = blockIdx.x;
    const int64_t row = tokens[m];
    for (int i = threadIdx.x; i < n; i += 256) {
        float v = QT == 2 ? packed_elem_6(w, row, i, nbp) : QT ? packed_elem_k(w, row, i, nbp) : packed_elem(w, row, i, nbp);
        if (scale != 1.0f) v = v * scale;
        x[(int64_t)m * n + i] = v;
    }
}
void launch_prefill_embed(int qt, const void *d_packed, int64_t cols, const int *d_tokens, int M, float scale, float *d_x, cudaStream_t st) {
    if (qt == 2) k_prefill_embed<2><<<M, 256, 0, st>>>((const uint8_t *)d_packed, nblocks_padded(cols), (int)cols, d_tokens, scale, d_x);
    else if (qt) k_prefill_embed<1><<<M, 256, 0, st>>>((const uint8_t *)d_packed, nblocks_padded(cols), (int)cols, d_tokens, scale, d_x);
    else k_prefill_embed<0><<<M, 256, 0, st>>>((const uint8_t *)d_packed, nblocks_padded(cols), (int)cols, d_tokens, scale, d_x);
}

// RoPE + KV append for M tokens at positions pos0..pos0+M-1: grid (nh + 2 nkv, M).  Same arithmetic as the decode kernels
// (oracle/model.py:_rope): q rotated in place in the QKV batch, k rotated and v rounded to f16 into the paged cache.
__global__ void __launch_bounds__(128) k_prefill_rope_kv(float *__restrict__ qkv, int64_t ld, int nh, int nkv, int hd, int pos0,
                                                         const float *__restrict__ cs, const float *__restrict__ sn, __half *__restrict__ kc,
                                                         __half *__restrict__ vc, const int32_t *__restrict__ page_table) {
    const int m = blockIdx.y, pos = pos0 + m, h = blockIdx.x, half = hd >> 1;
    const int page = page_table[pos / PAGE_TOKENS], slot = pos % PAGE_TOKENS;
    float *base = qkv + (int64_t)m * ld;
    if (h < nh + nkv) {
        float *t = base + (int64_t)h * hd;   // q heads then k heads are contiguous in the fused QKV output
        for (int j = threadIdx.x; j < half; j += blockDim.x) {
            const float c = cs[(int64_t)pos * half + j], s = sn[(int64_t)pos * half + j];
            const float a = t[j], b = t[j + half];
            const float ra = __fsub_rn(__fmul_rn(a, c), __fmul_rn(b, s));
            const float rb = __fadd_rn(__fmul_rn(b, c), __fmul_rn(a, s));
            if (h < nh) {
                t[j] = ra;
                t[j + half] = rb;
            } else {
                __half *dst = kc + (((int64_t)page * nkv + (h - nh)) * PAGE_TOKENS + slot) * hd;
                dst[j] = __float2half_rn(ra);
                dst[j + half] = __float2half_rn(rb);
            }
        }
    } else {
        const int kvh = h - nh - nkv;
        const float *t = base + (int64_t)(nh + nkv + kvh) * hd;
        __half *dst = vc + (((int64_t)page * nkv + kvh) * PAGE_TOKENS + slot) * hd;
        for (int j = threadIdx.x; j < hd; j += blockDim.x) dst[j] = __float2half_rn(t[j]);
    }
}

// Causal attention of a prompt batch over the paged f16 cache: grid (nh, M), query (head h, token m) sees positions
// [0, pos0 + m].  One CTA per query, scores in shared memory (max_ctx floats), fp32 softmax as in the decode kernel.
__global__ void __launch_bounds__(128) k_prefill_attn(const float *__restrict__ qkv, int64_t ld, int nh, int nkv, int hd, int pos0,
                                                      const __half *__restrict__ kc, const __half *__restrict__ vc,
                                                      const int32_t *__restrict__ page_table, float *__restrict__ out, int64_t ldo) {
    extern __shared__ float psm[];
    float *sq = psm;          // hd
    float *sc = psm + hd;     // n_ctx scores
    __shared__ float red[4];
    const int h = blockIdx.x, m = blockIdx.y, kvh = h / (nh / nkv);
    const int n_ctx = pos0 + m + 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_pages = (n_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const float scale = 1.0f / sqrtf((float)hd);
    for (int j = tid; j < hd; j += 128) sq[j] = qkv[(int64_t)m * ld + (int64_t)h * hd + j];
    __syncthreads();
    const int key = tid >> 3, sub = tid & 7, nchunk = hd >> 3;
    float lmax = -INFINITY;
    for (int p = 0; p < n_pages; ++p) {
        const int t = p * PAGE_TOKENS + key;
        const __half *kr = kc + (((int64_t)page_table[p] * nkv + kvh) * PAGE_TOKENS + key) * hd;
        float dot = 0.f;
        if (t < n_ctx) {
            for (int c = sub; c < nchunk; c += 8) {
                const uint4 raw = *(const uint4 *)(kr + c * 8);
                const __half2 *h2 = (const __half2 *)&raw;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 f = __half22float2(h2[e]);
                    dot = fmaf(sq[c * 8 + 2 * e], f.x, dot);
                    dot = fmaf(sq[c * 8 + 2 * e + 1], f.y, dot);
                }
            }
        }
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        dot += __shfl_xor_sync(0xffffffffu, dot, 4);
        if (t < n_ctx) {
            const float s = dot * scale;
            if (sub == 0) sc[t] = s;
            lmax = fmaxf(lmax, s);
        }
    }
    lmax = warp_max(lmax);
    if (lane == 0) red[warp] = lmax;
    __syncthreads();
    const float mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    __syncthreads();
    float lsum = 0.f;
    for (int i = tid; i < n_ctx; i += 128) {
        const float pe = expf(sc[i] - mx);
        sc[i] = pe;
        lsum += pe;
    }
    lsum = warp_sum(lsum);
    if (lane == 0) red[warp] = lsum;
    __syncthreads();
    const float l = red[0] + red[1] + red[2] + red[3];
    float o0 = 0.f, o1 = 0.f;
    const int d0 = tid, d1 = tid + 128;
    for (int p = 0; p < n_pages; ++p) {
        const __half *vr = vc + ((int64_t)page_table[p] * nkv + kvh) * PAGE_TOKENS * hd;
        const int tb = p * PAGE_TOKENS, lim = min(PAGE_TOKENS, n_ctx - tb);
        for (int k = 0; k < lim; ++k) {
            const float pe = sc[tb + k];
            if (d0 < hd) o0 = fmaf(pe, __half2float(vr[k * hd + d0]), o0);
            if (d1 < hd) o1 = fmaf(pe, __half2float(vr[k * hd + d1]), o1);
        }
    }
    if (d0 < hd) out[(int64_t)m * ldo + (int64_t)h * hd + d0] = o0 / l;
    if (d1 < hd) out[(int64_t)m * ldo + (int64_t)h * hd + d1] = o1 / l;
}

// Tiled form for whole prompt batches: one CTA per (head, block of 16 queries); a page of 16 keys / values is staged in shared memory
// once for the 16 queries (the one-CTA-per-query kernel above re-reads the cache for every query: 2 000-token prompts spent half
// their time there).  Thread (query qi = tid / 8, lane l8 = tid % 8) owns DPL = hd / 8 head dimensions: partial dot products over
// its dimensions for the 16 keys, three shuffles per key to complete them, online softmax (running maximum and sum per query),
// and the weighted values of its own dimensions.  fp32 throughout, f16 K / V as stored.  Head dim 256 (gemma) runs LPQ = 16 lanes per
// query (256 threads, four shuffles), which keeps DPL = 16 dimensions per thread: with 8 lanes and 32 dimensions per thread the kernel
// sat at 128 registers and lost to the one-CTA-per-query kernel.
template <int DPL, int LPQ>
__global__ void __launch_bounds__(16 * LPQ) k_prefill_attn_tiled(const float *__restrict__ qkv, int64_t ld, int nh, int nkv, int M, int pos0,
                                                            const __half *__restrict__ kc, const __half *__restrict__ vc,
                                                            const int32_t *__restrict__ page_table, float *__restrict__ out, int64_t ldo) {
    constexpr int HD = DPL * LPQ, NT = 16 * LPQ;
    __shared__ __align__(16) __half sk[PAGE_TOKENS * HD], sv[PAGE_TOKENS * HD];
    const int h = blockIdx.x, qb = blockIdx.y, kvh = h / (nh / nkv);
    const int tid = threadIdx.x, qi = tid / LPQ, l8 = tid % LPQ;
    const int m = qb * 16 + qi;                       // query token of this thread (may be >= M in the last block)
    const int qpos = pos0 + min(m, M - 1);
    const float scale = 1.0f / sqrtf((float)HD);
    float q[DPL], o[DPL];
#pragma unroll
    for (int d = 0; d < DPL; ++d) {
        q[d] = qkv[(int64_t)min(m, M - 1) * ld + (int64_t)h * HD + l8 * DPL + d] * scale;
        o[d] = 0.f;
    }
    float mx = -INFINITY, sum = 0.f;
    const int last_pos = pos0 + min(qb * 16 + 15, M - 1);          // the newest position any query of this block may see
    const int n_pages = last_pos / PAGE_TOKENS + 1;
    // the next page travels L2 -> registers while the current one is worked on (a CTA walks its pages one after the other; without
    // this every page costs an exposed L2 round trip)
    constexpr int NVU = (PAGE_TOKENS * HD / 8 + NT - 1) / NT;       // uint4 per thread and matrix (hd 96: 1.5 -> 2 with a bound check)
    uint4 rk[NVU], rv[NVU];
    auto fetch = [&](int p) {
        const size_t base = ((size_t)page_table[p] * nkv + kvh) * PAGE_TOKENS * HD;
#pragma unroll
        for (int u = 0; u < NVU; ++u) {
            const int i = tid + u * NT;
            rk[u] = rv[u] = make_uint4(0u, 0u, 0u, 0u);
            if (i < PAGE_TOKENS * HD / 8) {
                rk[u] = ((const uint4 *)(kc + base))[i];
                rv[u] = ((const uint4 *)(vc + base))[i];
            }
        }
    };
    fetch(0);
    for (int p = 0; p < n_pages; ++p) {
        __syncthreads();                                           // the previous page is consumed
#pragma unroll
        for (int u = 0; u < NVU; ++u) {
            const int i = tid + u * NT;
            if (i < PAGE_TOKENS * HD / 8) { ((uint4 *)sk)[i] = rk[u]; ((uint4 *)sv)[i] = rv[u]; }
        }
        if (p + 1 < n_pages) fetch(p + 1);
        __syncthreads();
        float sc[PAGE_TOKENS];
#pragma unroll
        for (int k = 0; k < PAGE_TOKENS; ++k) {
            float dot = 0.f;
#pragma unroll
            for (int d = 0; d < DPL; d += 2) {
                const float2 kk = __half22float2(*(const __half2 *)(sk + k * HD + l8 * DPL + d));
                dot = fmaf(q[d], kk.x, dot);
                dot = fmaf(q[d + 1], kk.y, dot);
            }
            dot += __shfl_xor_sync(0xffffffffu, dot, 1);
            dot += __shfl_xor_sync(0xffffffffu, dot, 2);
            dot += __shfl_xor_sync(0xffffffffu, dot, 4);
            if (LPQ == 16) dot += __shfl_xor_sync(0xffffffffu, dot, 8);
            sc[k] = (p * PAGE_TOKENS + k <= qpos) ? dot : -INFINITY;   // causal mask (and slots not written yet)
        }
        float pm = mx;
#pragma unroll
        for (int k = 0; k < PAGE_TOKENS; ++k) pm = fmaxf(pm, sc[k]);
        if (pm == -INFINITY) continue;                             // nothing visible in this page for this query (uniform per query group)
        const float corr = expf(mx - pm);                          // exp(-inf) = 0 on the first visible page
        sum *= corr;
#pragma unroll
        for (int d = 0; d < DPL; ++d) o[d] *= corr;
#pragma unroll
        for (int k = 0; k < PAGE_TOKENS; ++k) {
            const float pe = expf(sc[k] - pm);
            sum += pe;
#pragma unroll
            for (int d = 0; d < DPL; d += 2) {
                const float2 vv = __half22float2(*(const __half2 *)(sv + k * HD + l8 * DPL + d));
                o[d] = fmaf(pe, vv.x, o[d]);
                o[d + 1] = fmaf(pe, vv.y, o[d + 1]);
            }
        }
        mx = pm;
    }
    if (m < M) {
#pragma unroll
        for (int d = 0; d < DPL; ++d) out[(int64_t)m * ldo + (int64_t)h * HD + l8 * DPL + d] = o[d] / sum;
    }
}

cudaError_t launch_prefill_attention(float *d_qkv, int64_t ld, int M, int nh, int nkv, int hd, int pos0, const float *d_cos, const float *d_sin,
                                     __half *d_k, __half *d_v, const int32_t *d_page_table, float *d_out, int64_t ldo, cudaStream_t st) {
    k_prefill_rope_kv<<<dim3(nh + 2 * nkv, M), 128, 0, st>>>(d_qkv, ld, nh, nkv, hd, pos0, d_cos, d_sin, d_k, d_v, d_page_table);
    const dim3 grid(nh, (M + 15) / 16);
    // short contexts: one CTA per query has 16x the CTAs and wins (32-token prompt 9.2 vs 9.6 ms); long ones: the tiled kernel reads
    // the cache once per 16 queries (2 000-token prompt, llama3.1:8b: 763 -> 517 ms).  hd 256 (gemma): 16 lanes per query; with 8 lanes
    // (32 dimensions per thread, 128 registers) the tiled kernel measured slower than one CTA per query (716 vs 615 ms).
    static int tiled256 = -1;   // B2_PF_ATTN_TILED256=0: head dim 256 back on the per-query kernel
    if (tiled256 < 0) { const char *e = getenv("B2_PF_ATTN_TILED256"); tiled256 = (e && e[0] == '0') ? 0 : 1; }
    const bool tiled = pos0 + M > 256;
    // head dim 128 with 16 lanes per query as well (8 dimensions per thread, twice the threads per CTA): llama3.1:8b 2 000-token prompt
    // 271 -> 261 ms, 512 tokens 53.6 -> 53.0; B2_PF_ATTN_LPQ16=0: 8 lanes
    static int lpq16 = -1;
    if (lpq16 < 0) { const char *e = getenv("B2_PF_ATTN_LPQ16"); lpq16 = (e && e[0] == '0') ? 0 : 1; }
    if (tiled && hd == 128 && lpq16) k_prefill_attn_tiled<8, 16><<<grid, 256, 0, st>>>(d_qkv, ld, nh, nkv, M, pos0, d_k, d_v, d_page_table, d_out, ldo);
    else if (tiled && hd == 128) k_prefill_attn_tiled<16, 8><<<grid, 128, 0, st>>>(d_qkv, ld, nh, nkv, M, pos0, d_k, d_v, d_page_table, d_out, ldo);
    else if (tiled && hd == 96) k_prefill_attn_tiled<12, 8><<<grid, 128, 0, st>>>(d_qkv, ld, nh, nkv, M, pos0, d_k, d_v, d_page_table, d_out, ldo);
    else if (tiled && hd == 256 && tiled256) k_prefill_attn_tiled<16, 16><<<grid, 256, 0, st>>>(d_qkv, ld, nh, nkv, M, pos0, d_k, d_v, d_page_table, d_out, ldo);
    else {   // one CTA per query
        const size_t smem = (size_t)(hd + pos0 + M) * sizeof(float);
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(k_prefill_attn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        k_prefill_attn<<<dim3(nh, M), 128, smem, st>>>(d_qkv, ld, nh, nkv, hd, pos0, d_k, d_v, d_page_table, d_out, ldo);
    }
    return cudaGetLastError();
}

// -inf / max-int in every arg-max partial (the LM-head GEMV after a prompt batch fills only its own grid's entries)
__global__ void k_fill_partials(float *pmax, int *pidx, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { pmax[i] = -INFINITY; pidx[i] = 0x7fffffff; }
}
void launch_fill_partials(float *pmax, int *pidx, int n, cudaStream_t st) { k_fill_partials<<<(n + 255) / 256, 256, 0, st>>>(pmax, pidx, n); }
