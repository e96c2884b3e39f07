// gemv.cu — the streaming Q4_0 GEMV of the decode path (stages a3.3 QKV, a3.6 O, a3.7 gate/up+act*mul,
// a3.8 down and LM head of Ollama's /api/generate loop; SURVEY.md §3.3, call site experiment/RunnerConfig.py:128-131).
//
// One persistent CTA of 16 warps per SM; the CTA owns a contiguous range of 16-row weight tiles.  A tile is
// cut along K into chunks of 64 blocks = 16 block-groups; warp w owns block-group w of every chunk, so its
// weight stream is a sequence of 1 KB (nibbles) + 128 B (scales) pieces that it loads straight into registers
// with 128-bit streaming loads (ld.global.nc.L1::no_allocate), GEMV_DEPTH pieces in flight.  The first loads are
// issued before griddepcontrol.wait: weights do not depend on the previous kernel.  The activation digits a
// warp needs (the B fragments of its block-groups) are held in registers, so the inner loop touches no shared
// memory.  Per block-group: rebuild int8 A fragments from the nibbles (two's-complement nibble in the high
// half of each byte = 16*(q-8): 3 ALU ops per 8 weights), four mma.sync.m16n8k32.s8.s8.s32 (16 rows x 32 weights
// x 3 digit planes of exact integer dot products each), scale by d_w * 2^e / 16 into fp32 accumulators.
// The sixteen warps split K and meet in shared memory once per tile.  Measured alone (tools/microbench3.cu)
// this structure streams 6.3-6.8 TB/s from HBM with the arithmetic included.
#include "gemv_launch.cuh"
#include "randinit.cuh"
#include <stdlib.h>

// ------------------------------------------------------------------------------------------ pack / init
// GGML Q4_0 row-major blocks (18 B: f16 d | 16 B nibbles) -> streaming layout (layout.cuh). One thread per block.
__global__ void k_pack_q4_0(const uint8_t *__restrict__ src, int64_t rows, int64_t nb, int64_t nbp, uint8_t *__restrict__ dst,
                            int64_t dst_row0, int64_t dst_row_stride, int64_t snb, int64_t sb0) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    int64_t r = i / nb, b = i - r * nb;
    const uint8_t *p = src + (r * snb + sb0 + b) * 18;
    PackedLoc loc = packed_loc(dst_row0 + r * dst_row_stride, b, nbp);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        uint32_t w = (uint32_t)p[2 + 4 * t] | ((uint32_t)p[3 + 4 * t] << 8) | ((uint32_t)p[4 + 4 * t] << 16) | ((uint32_t)p[5 + 4 * t] << 24);
        *(uint32_t *)(dst + loc.word0 + t * 16) = w ^ 0x88888888u;
    }
    *(uint16_t *)(dst + loc.scale) = (uint16_t)p[0] | ((uint16_t)p[1] << 8);
}
void launch_pack_q4_0(const void *d_src, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                      int64_t src_cols, int64_t src_col0) {
    int64_t nb = cols / QK, n = rows * nb;
    const int64_t snb = src_cols ? src_cols / QK : nb, sb0 = src_col0 / QK;   // the source rows are snb blocks long; take blocks [sb0, sb0 + nb)
    k_pack_q4_0<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const uint8_t *)d_src, rows, nb, nblocks_padded(cols), (uint8_t *)d_packed, dst_row0, dst_row_stride, snb, sb0);
}
// GGML Q4_K row-major super-blocks (144 B: f16 d | f16 dmin | 12 B packed 6-bit scales/mins | 128 B nibbles) -> streaming
// layout (layout.cuh).  One thread per 32-element sub-block.  ggml: sub-blocks 2p and 2p+1 share the bytes qs[32p..32p+31]
// (low / high nibbles); scale and min of sub-block j follow get_scale_min_k4.
__global__ void k_pack_q4_k(const uint8_t *__restrict__ src, int64_t rows, int64_t nb, int64_t nbp, uint8_t *__restrict__ dst,
                            int64_t dst_row0, int64_t dst_row_stride, int64_t snb, int64_t sb0) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    const int64_t r = i / nb, b = i - r * nb;
    const int j = (int)(b & 7);   // sb0 is a multiple of 8: a column window starts on a super-block
    const uint8_t *p = src + (r * (snb >> 3) + ((sb0 + b) >> 3)) * 144;
    const uint8_t *q = p + 4, *qs = p + 16 + 32 * (j >> 1);
    uint8_t sc, m;
    if (j < 4) { sc = q[j] & 63; m = q[j + 4] & 63; }
    else { sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
    const int sh = 4 * (j & 1);
    PackedLocK loc = packed_loc_k(dst_row0 + r * dst_row_stride, b, nbp);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        uint32_t w = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int e = 4 * t + k;
            const uint32_t lo = (qs[e] >> sh) & 0xF, hi = (qs[e + 16] >> sh) & 0xF;
            w |= (lo | (hi << 4)) << (8 * k);
        }
        *(uint32_t *)(dst + loc.word0 + t * 16) = w ^ 0x88888888u;
    }
    dst[loc.scm] = sc;
    dst[loc.scm + 1] = m;
    *(uint32_t *)(dst + loc.dm) = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);   // every sub-block writes the same value
}
void launch_pack_q4_k(const void *d_src, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                      int64_t src_cols, int64_t src_col0) {
    int64_t nb = cols / QK, n = rows * nb;
    const int64_t snb = src_cols ? src_cols / QK : nb, sb0 = src_col0 / QK;   // the source rows are snb blocks long; take blocks [sb0, sb0 + nb)
    k_pack_q4_k<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const uint8_t *)d_src, rows, nb, nblocks_padded(cols), (uint8_t *)d_packed, dst_row0, dst_row_stride, snb, sb0);
}
// GGML Q6_K row-major super-blocks (210 B: ql[128] | qh[64] | int8 scales[16] | f16 d) -> streaming layout (layout.cuh).
// One thread per 32-element block jb of a super-block: jb = 4n + k covers elements l = 0..31 of ggml's group (n, k):
// low nibble from ql[64n + 32(k&1) + l] (low half for k < 2, high half otherwise), high bits from (qh[32n + l] >> 2k) & 3.
__global__ void k_pack_q6_k(const uint8_t *__restrict__ src, int64_t rows, int64_t nb, int64_t nbp, uint8_t *__restrict__ dst,
                            int64_t dst_row0, int64_t dst_row_stride, int64_t snb, int64_t sb0) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    const int64_t r = i / nb, b = i - r * nb;
    const int jb = (int)(b & 7), n = jb >> 2, k = jb & 3;
    const uint8_t *p = src + (r * (snb >> 3) + ((sb0 + b) >> 3)) * 210;
    const uint8_t *ql = p + 64 * n + 32 * (k & 1), *qh = p + 128 + 32 * n;
    const int lsh = (k >> 1) * 4, hsh = 2 * k;
    PackedLoc6 loc = packed_loc_6(dst_row0 + r * dst_row_stride, b, nbp);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        uint32_t w = 0, hb = 0;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const int e = 4 * t + kk;
            const uint32_t lo = (ql[e] >> lsh) & 0xF, lo2 = (ql[e + 16] >> lsh) & 0xF;
            w |= (lo | (lo2 << 4)) << (8 * kk);
            hb |= (((uint32_t)qh[e] >> hsh) & 3u) << (2 * kk);
            hb |= (((uint32_t)qh[e + 16] >> hsh) & 3u) << (8 + 2 * kk);
        }
        *(uint32_t *)(dst + loc.word0 + t * 16) = w ^ 0x88888888u;
        *(uint16_t *)(dst + loc.hi + t * 16) = (uint16_t)hb;
    }
    dst[loc.sc] = p[192 + 2 * jb];
    dst[loc.sc + 1] = p[192 + 2 * jb + 1];
    *(uint16_t *)(dst + loc.d) = (uint16_t)p[208] | ((uint16_t)p[209] << 8);
}
void launch_pack_q6_k(const void *d_src, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                      int64_t src_cols, int64_t src_col0) {
    int64_t nb = cols / QK, n = rows * nb;
    const int64_t snb = src_cols ? src_cols / QK : nb, sb0 = src_col0 / QK;   // the source rows are snb blocks long; take blocks [sb0, sb0 + nb)
    k_pack_q6_k<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const uint8_t *)d_src, rows, nb, nblocks_padded(cols), (uint8_t *)d_packed, dst_row0, dst_row_stride, snb, sb0);
}
// Deterministic random init straight into the streaming layout; same bytes as b2_randinit_tensor + pack.
__global__ void k_randinit_q4_0(uint64_t key, float sigma, int64_t rows, int64_t nb, int64_t nbp, uint8_t *__restrict__ dst,
                                int64_t dst_row0, int64_t dst_row_stride, int64_t sr0, int64_t snb, int64_t sb0) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    int64_t r = i / nb, b = i - r * nb;
    float x[QK];
    ri_gauss_block(key, (uint64_t)((sr0 + r) * snb + sb0 + b) * QK, sigma, x);   // block (sr0 + r, sb0 + b) of the whole tensor
    uint16_t dbits;
    uint32_t w[4];
    ri_quant_q4_0(x, dbits, w);
    PackedLoc loc = packed_loc(dst_row0 + r * dst_row_stride, b, nbp);
#pragma unroll
    for (int t = 0; t < 4; ++t) *(uint32_t *)(dst + loc.word0 + t * 16) = w[t] ^ 0x88888888u;
    *(uint16_t *)(dst + loc.scale) = dbits;
}
void launch_randinit_q4_0(uint64_t key, float sigma, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                          int64_t src_row0, int64_t src_cols, int64_t src_col0) {
    int64_t nb = cols / QK, n = rows * nb;
    k_randinit_q4_0<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(key, sigma, rows, nb, nblocks_padded(cols), (uint8_t *)d_packed, dst_row0, dst_row_stride,
                                                                 src_row0, src_cols ? src_cols / QK : nb, src_col0 / QK);
}
// Synthetic Q4_K blocks straight into the streaming layout (benchmarks of SURVEY.md 8d config 4; no checkpoint exists offline
// and ggml's Q4_K search quantiser is not worth running over 8 G weights): per super-block d = f16(9e-5 * (0.75 + u/2)),
// dmin = 8 d; per sub-block sc in [32, 63], m = round(7.5 sc / 8) so that d*sc*q - dmin*m is centred; nibbles from the hash.
// The dequantised weights have mean ~0 and standard deviation ~0.02 like the Q4_0 initialiser's.
__global__ void k_randinit_q4_k(uint64_t key, int64_t rows, int64_t nb, int64_t nbp, uint8_t *__restrict__ dst, int64_t dst_row0, int64_t dst_row_stride,
                                int64_t sr0, int64_t snb, int64_t sb0) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    const int64_t r = i / nb, b = i - r * nb;
    const int64_t fr = sr0 + r, fb = sb0 + b;   // coordinates in the whole tensor (a rank of a tensor-parallel model fills its window)
    i = fr * snb + fb;
    const uint64_t hs = ri_splitmix64(key ^ (0x5bd1e995ull * (uint64_t)(fr * (snb >> 3) + (fb >> 3) + 1)));   // per super-block
    const uint64_t h0 = ri_splitmix64(key + 2 * (uint64_t)i), h1 = ri_splitmix64(key + 2 * (uint64_t)i + 1);
    const float d = 9e-5f * (0.75f + 0.5f * (float)(hs & 0xFFFF) * (1.0f / 65536.0f));
    const __half dh = __float2half_rn(d), dmh = __float2half_rn(8.0f * __half2float(dh));
    const uint32_t sc = 32u + (uint32_t)((hs >> (16 + 5 * (fb & 7))) & 31u), mn = (15u * sc + 8u) >> 4;
    PackedLocK loc = packed_loc_k(dst_row0 + r * dst_row_stride, b, nbp);
    *(uint32_t *)(dst + loc.word0) = (uint32_t)h0;
    *(uint32_t *)(dst + loc.word0 + 16) = (uint32_t)(h0 >> 32);
    *(uint32_t *)(dst + loc.word0 + 32) = (uint32_t)h1;
    *(uint32_t *)(dst + loc.word0 + 48) = (uint32_t)(h1 >> 32);
    dst[loc.scm] = (uint8_t)sc;
    dst[loc.scm + 1] = (uint8_t)mn;
    *(__half2 *)(dst + loc.dm) = __halves2half2(dh, dmh);
}
void launch_randinit_q4_k(uint64_t key, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                          int64_t src_row0, int64_t src_cols, int64_t src_col0) {
    int64_t nb = cols / QK, n = rows * nb;
    k_randinit_q4_k<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(key, rows, nb, nblocks_padded(cols), (uint8_t *)d_packed, dst_row0, dst_row_stride,
                                                                 src_row0, src_cols ? src_cols / QK : nb, src_col0 / QK);
}
// Synthetic Q6_K blocks (benchmarks with a Q6_K LM head, as in llama.cpp's "Q4_0" files): random 6-bit values, int8 scales in
// [32, 127], d ~ 1.35e-5 so that d*sc*(q-32) has standard deviation ~0.02.
__global__ void k_randinit_q6_k(uint64_t key, int64_t rows, int64_t nb, int64_t nbp, uint8_t *__restrict__ dst, int64_t dst_row0, int64_t dst_row_stride,
                                int64_t sr0, int64_t snb, int64_t sb0) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    const int64_t r = i / nb, b = i - r * nb;
    const int64_t fr = sr0 + r, fb = sb0 + b;   // coordinates in the whole tensor (a rank of a tensor-parallel model fills its window)
    i = fr * snb + fb;
    const uint64_t hs = ri_splitmix64(key ^ (0x5bd1e995ull * (uint64_t)(fr * (snb >> 3) + (fb >> 3) + 1)));   // per super-block
    const uint64_t h0 = ri_splitmix64(key + 3 * (uint64_t)i), h1 = ri_splitmix64(key + 3 * (uint64_t)i + 1), h2 = ri_splitmix64(key + 3 * (uint64_t)i + 2);
    const float d = 1.35e-5f * (0.75f + 0.5f * (float)(hs & 0xFFFF) * (1.0f / 65536.0f));
    PackedLoc6 loc = packed_loc_6(dst_row0 + r * dst_row_stride, b, nbp);
    *(uint32_t *)(dst + loc.word0) = (uint32_t)h0;
    *(uint32_t *)(dst + loc.word0 + 16) = (uint32_t)(h0 >> 32);
    *(uint32_t *)(dst + loc.word0 + 32) = (uint32_t)h1;
    *(uint32_t *)(dst + loc.word0 + 48) = (uint32_t)(h1 >> 32);
#pragma unroll
    for (int t = 0; t < 4; ++t) *(uint16_t *)(dst + loc.hi + t * 16) = (uint16_t)(h2 >> (16 * t));
    dst[loc.sc] = (uint8_t)(32u + (uint32_t)((hs >> 16) & 63u) + (uint32_t)((hs >> (22 + 4 * (fb & 7))) & 31u));
    dst[loc.sc + 1] = (uint8_t)(32u + (uint32_t)((hs >> 19) & 63u) + (uint32_t)((hs >> (24 + 4 * (fb & 7))) & 31u));
    *(__half *)(dst + loc.d) = __float2half_rn(d);
}
void launch_randinit_q6_k(uint64_t key, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                          int64_t src_row0, int64_t src_cols, int64_t src_col0) {
    int64_t nb = cols / QK, n = rows * nb;
    k_randinit_q6_k<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(key, rows, nb, nblocks_padded(cols), (uint8_t *)d_packed, dst_row0, dst_row_stride,
                                                                 src_row0, src_cols ? src_cols / QK : nb, src_col0 / QK);
}
__global__ void k_randinit_f32(uint64_t key, float sigma, float offset, int64_t n, float *__restrict__ out, int64_t i0) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ri_add(ri_mul(ri_gauss(key, (uint64_t)(i0 + i)), sigma), offset);
}
void launch_randinit_f32(uint64_t key, float sigma, float offset, int64_t n, float *d_out, cudaStream_t st, int64_t i0) {
    k_randinit_f32<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(key, sigma, offset, n, d_out, i0);
}
// ------------------------------------------------------------------------------------------ launcher
static int g_num_sms = 0;
static long long *g_gemv_dbg = nullptr;   // set by b2dbg_gemv_profile (development aid): ring of `slots` x [148][8] stamps
static int g_gemv_dbg_slots = 1, g_gemv_dbg_next = 0;
extern "C" void b2dbg_gemv_profile(long long *d_buf) { g_gemv_dbg = d_buf; g_gemv_dbg_slots = 1; }
extern "C" void b2dbg_gemv_profile_ring(long long *d_buf, int slots) { g_gemv_dbg = d_buf; g_gemv_dbg_slots = slots; g_gemv_dbg_next = 0; }
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

// Fill the argument block of one GEMV (everything but the debug pointer) and report the dynamic shared memory it needs.
cudaError_t gemv_make_args(GemvArgs &a, int in_mode, const void *d_packed, int64_t rows, int64_t cols, const float *x, const float *norm_w,
                           float eps, int plus_one, const void *act_in, float *y, const float *bias_or_res, void *act_out, int act,
                           float *pmax, int *pidx, const void *next_w, uint64_t next_bytes, const float *xsq_in, int xsq_n, float *xsq_out,
                           size_t *smem_out, int qt) {
    a = GemvArgs{};
    a.qt = qt;
    a.xsq_in = xsq_in; a.xsq_n = xsq_n; a.xsq_out = xsq_out;
    a.w = (const uint8_t *)d_packed;
    a.rows = (int)rows;
    a.nbp = (int)nblocks_padded(cols);
    a.n_tiles = (int)ntiles(rows);
    a.tile_align = 1;
    if (in_mode == IN_NORM && (!xsq_in || xsq_n <= 0)) return cudaErrorInvalidValue;
    a.x = x; a.norm_w = norm_w; a.eps = eps; a.plus_one = plus_one; a.n_in = (int)cols;
    a.act_in = (const uint8_t *)act_in;
    a.y = y; a.bias_or_res = bias_or_res; a.act_out = (uint8_t *)act_out; a.n_out_act = (int)(rows / 2); a.act = act;
    a.pmax = pmax; a.pidx = pidx;
    a.pf_ptr = (const uint8_t *)next_w; a.pf_total = next_bytes;
    static int env_win = -1, env_pf = -1;   // tuning knobs: B2_L2_WINDOW / B2_PF_NEXT (bytes)
    if (env_win < 0) {
        const char *e1 = getenv("B2_L2_WINDOW"), *e2 = getenv("B2_PF_NEXT");
        env_win = e1 ? atoi(e1) : L2_WINDOW_BYTES;
        env_pf = e2 ? atoi(e2) : L2_PREFETCH_BYTES;
    }
    a.l2_window = (uint32_t)env_win;
    a.pf_bytes = (uint32_t)env_pf;
    static int env_early = -1;   // B2_PF_EARLY (bytes per CTA): the O projection, parked behind attention, pulls the head of the gate/up matrix into L2 before it waits
    if (env_early < 0) { const char *e3 = getenv("B2_PF_EARLY"); env_early = e3 ? atoi(e3) : L2_PREFETCH_EARLY_BYTES; }
    if (in_mode == IN_ACT) a.pf_bytes = (uint32_t)env_early;
    a.k_outer = (a.nbp + KC - 1) / KC > 2;
    // a flush costs two CTA barriers and stalls all sixteen streams: park as many tiles as there are warps to finalise them
    // (LM head 62.5 -> 60.3 us, +1.7 % tokens/s against 8); with chunks outermost a CTA has at most GEMV_MAXT tiles
    a.rb = a.k_outer ? GEMV_MAXT : GEMV_RB_MAX;
    const size_t smem = GEMV_FIXED_SMEM(a.rb) + (size_t)a.nbp * (100 + 4 * qt) + (in_mode == IN_ACT ? 0 : (size_t)cols * 4 + 1024) + 128;
    if (smem > 227 * 1024) return cudaErrorInvalidValue;
    if (smem_out) *smem_out = smem;
    return cudaSuccess;
}

// in_mode IN_NORM: (x, norm_w, eps, plus_one) ; IN_ACT: act_in.  out_mode selects the epilogue.
// next_w / next_bytes: the packed matrix the following kernel streams (L2 prefetch hint), or nullptr.
cudaError_t launch_gemv(int in_mode, int out_mode, const void *d_packed, int64_t rows, int64_t cols, const float *x, const float *norm_w,
                        float eps, int plus_one, const void *act_in, float *y, const float *bias_or_res, void *act_out, int act,
                        float *pmax, int *pidx, int *grid_out, const void *next_w, uint64_t next_bytes, const float *xsq_in, int xsq_n,
                        float *xsq_out, cudaStream_t st, bool pdl) {
    return launch_gemv_qt(0, in_mode, out_mode, d_packed, rows, cols, x, norm_w, eps, plus_one, act_in, y, bias_or_res, act_out, act, pmax, pidx,
                          grid_out, next_w, next_bytes, xsq_in, xsq_n, xsq_out, st, pdl);
}
// qt: 0 = Q4_0 matrix, 1 = Q4_K, 2 = Q6_K (layout.cuh)
cudaError_t launch_gemv_qt(int qt, int in_mode, int out_mode, const void *d_packed, int64_t rows, int64_t cols, const float *x, const float *norm_w,
                           float eps, int plus_one, const void *act_in, float *y, const float *bias_or_res, void *act_out, int act,
                           float *pmax, int *pidx, int *grid_out, const void *next_w, uint64_t next_bytes, const float *xsq_in, int xsq_n,
                           float *xsq_out, cudaStream_t st, bool pdl) {
    GemvArgs a;
    size_t smem = 0;
    cudaError_t em = gemv_make_args(a, in_mode, d_packed, rows, cols, x, norm_w, eps, plus_one, act_in, y, bias_or_res, act_out, act, pmax, pidx,
                                    next_w, next_bytes, xsq_in, xsq_n, xsq_out, &smem, qt);
    if (em != cudaSuccess) return em;
    a.dbg = g_gemv_dbg ? g_gemv_dbg + (size_t)(g_gemv_dbg_next++ % g_gemv_dbg_slots) * 148 * 8 : nullptr;
    int grid = gemv_grid(a.n_tiles, a.tile_align);
    const int units = (a.n_tiles + a.tile_align - 1) / a.tile_align;
    int max_tiles = ((units + grid - 1) / grid) * a.tile_align;
    if (grid_out) *grid_out = grid;
#define CASE(I, O) if (in_mode == I && out_mode == O) return launch_one<I, O>(aa, grid, smem, st, pdl)
    if (!a.k_outer || max_tiles <= GEMV_MAXT) {
        const GemvArgs &aa = a;
        CASE(IN_NORM, OUT_BIAS); CASE(IN_NORM, OUT_RES); CASE(IN_NORM, OUT_ACTMUL); CASE(IN_NORM, OUT_LOGITS);
        CASE(IN_ACT, OUT_BIAS); CASE(IN_ACT, OUT_RES); CASE(IN_ACT, OUT_ACTMUL);
        CASE(IN_F32, OUT_BIAS); CASE(IN_F32, OUT_RES);
        return cudaErrorInvalidValue;
    }
    // long K and many rows (no matrix of the seven models; kept for generality): the accumulators of a CTA's tiles live in
    // registers, so cover the rows with several launches of at most GEMV_MAXT tiles per CTA
    if (out_mode == OUT_LOGITS || out_mode == OUT_ACTMUL) return cudaErrorInvalidValue;
    const int per_launch = g_num_sms * GEMV_MAXT;
    for (int tbeg = 0; tbeg < a.n_tiles; tbeg += per_launch) {
        GemvArgs aa = a;
        const int tn = a.n_tiles - tbeg < per_launch ? a.n_tiles - tbeg : per_launch;
        aa.w = a.w + (uint64_t)tbeg * a.nbp * (a.qt == 2 ? BLOCK_BYTES_6 : a.qt ? BLOCK_BYTES_K : BLOCK_BYTES);
        aa.n_tiles = tn;
        aa.rows = (int)(rows - (int64_t)tbeg * TILE_ROWS < (int64_t)tn * TILE_ROWS ? rows - (int64_t)tbeg * TILE_ROWS : (int64_t)tn * TILE_ROWS);
        aa.y = y + (int64_t)tbeg * TILE_ROWS;
        aa.bias_or_res = bias_or_res ? bias_or_res + (int64_t)tbeg * TILE_ROWS : nullptr;
        aa.xsq_out = nullptr;
        grid = gemv_grid(tn, 1);
        cudaError_t e = cudaErrorInvalidValue;
        if (in_mode == IN_NORM && out_mode == OUT_BIAS) e = launch_one<IN_NORM, OUT_BIAS>(aa, grid, smem, st, pdl);
        if (in_mode == IN_NORM && out_mode == OUT_RES) e = launch_one<IN_NORM, OUT_RES>(aa, grid, smem, st, pdl);
        if (in_mode == IN_ACT && out_mode == OUT_BIAS) e = launch_one<IN_ACT, OUT_BIAS>(aa, grid, smem, st, pdl);
        if (in_mode == IN_ACT && out_mode == OUT_RES) e = launch_one<IN_ACT, OUT_RES>(aa, grid, smem, st, pdl);
        if (in_mode == IN_F32 && out_mode == OUT_BIAS) e = launch_one<IN_F32, OUT_BIAS>(aa, grid, smem, st, pdl);
        if (in_mode == IN_F32 && out_mode == OUT_RES) e = launch_one<IN_F32, OUT_RES>(aa, grid, smem, st, pdl);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
#undef CASE
}

// ------------------------------------------------------------------------------------------ embedding row
// x = embedding row of the current token (gemma: * sqrt(H)); also leaves sum(x^2) for the first RMSNorm.  One block.
template <int QT>
__global__ void __launch_bounds__(1024) k_embed_row(const uint8_t *__restrict__ w, int64_t nbp, int n, const int *__restrict__ token, float scale,
                                                    float *__restrict__ x, float *__restrict__ xsq) {
    __shared__ float red[32];
    const int64_t row = __ldcg(token);
    float ss = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float v = QT == 2 ? packed_elem_6(w, row, i, nbp) : QT ? packed_elem_k(w, row, i, nbp) : packed_elem(w, row, i, nbp);
        if (scale != 1.0f) v = v * scale;
        x[i] = v;
        ss = fmaf(v, v, ss);
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x < 32) {
        float tv = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
        tv = warp_sum(tv);
        if (threadIdx.x == 0) xsq[0] = tv;
    }
}
void launch_embed_row(int qt, const void *d_packed, int64_t cols, const int *d_token, float scale, float *d_x, float *d_xsq, cudaStream_t st) {
    if (qt == 2) k_embed_row<2><<<1, 1024, 0, st>>>((const uint8_t *)d_packed, nblocks_padded(cols), (int)cols, d_token, scale, d_x, d_xsq);
    else if (qt) k_embed_row<1><<<1, 1024, 0, st>>>((const uint8_t *)d_packed, nblocks_padded(cols), (int)cols, d_token, scale, d_x, d_xsq);
    else k_embed_row<0><<<1, 1024, 0, st>>>((const uint8_t *)d_packed, nblocks_padded(cols), (int)cols, d_token, scale, d_x, d_xsq);
}
// sum of squares of an f32 vector into xsq[0] (op-level entry points that run the fused RMSNorm on a caller's vector)
__global__ void __launch_bounds__(1024) k_sumsq(const float *__restrict__ x, int n, float *__restrict__ xsq) {
    __shared__ float red[32];
    float ss = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { const float v = x[i]; ss = fmaf(v, v, ss); }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x < 32) {
        float tv = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
        tv = warp_sum(tv);
        if (threadIdx.x == 0) xsq[0] = tv;
    }
}
void launch_sumsq(const float *d_x, int64_t n, float *d_xsq, cudaStream_t st) { k_sumsq<<<1, 1024, 0, st>>>(d_x, (int)n, d_xsq); }
