// gemv_dev.cuh — device side of the streaming Q4_0 GEMV (see gemv.cu for the design): shared by the single-GPU kernels
// (gemv.cu) and the tensor-parallel ones (gemv_tp.cu).
#pragma once
#include "kernels.cuh"
#include "layout.cuh"

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
__device__ __forceinline__ bool mbar_test(uint64_t *b, uint32_t parity) {   // non-blocking
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
#ifdef GEMV_EXP_NOWAIT
__device__ __forceinline__ void pdl_wait() {}   // what-if experiment: no ordering against the previous kernel (results are garbage)
#else
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void mma_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// ------------------------------------------------------------------------------------------ the kernel
#define GEMV_CWARPS 16                       // one block-group of a 64-block chunk each
#define GEMV_THREADS (GEMV_CWARPS * 32)
#define GEMV_DEPTH 4                         // weight pieces (1152 B per warp) in flight per warp
#define GEMV_NORM_MAX 8192                   // longest vector the fused RMSNorm prologue keeps in registers
#define GEMV_MAXT 2                          // K-outer order: tiles per CTA whose accumulators stay in registers
#define MMA_MAGIC 0x4B400000                 // int bits of 12582912.0f: accumulators start here, so c reinterpreted as float is 12582912 + S exactly
#define MMA_MAGIC_F 12582912.0f
#define L2_PREFETCH_BYTES (64 * 1024)        // of the NEXT kernel's weights, per CTA, issued behind this kernel's own stream (B2_PF_NEXT): the next kernel's first
                                             // pieces then come from L2 instead of HBM (656.9 -> 675 tokens/s at 128 KB alone, 687 together with the early prefetch)
#define L2_PREFETCH_EARLY_BYTES (192 * 1024) // the O projection, resident and parked behind attention, pulls this much of the gate/up matrix per CTA into L2 BEFORE
                                             // its dependency wait (B2_PF_EARLY; 28 MB in all): HBM works through the attention (128 / 192 / 256 / 320 / 384 KB:
                                             // 667 / 683 / 678-680 / 674 / 668 tokens/s against 657)
#define L2_WINDOW_BYTES (128 * 1024)         // how far the CTA's L2 prefetch cursor runs ahead of its register loads
#ifndef GEMV_ORDER
#define GEMV_ORDER 0                         // stand-alone kernel: 0 = register loads, wait, staging; 1 = wait, staging, register loads; 2 = as 0 with the L2 window after the wait
#endif
#ifndef MMA_GROUP
#define MMA_GROUP 2                          // independent MMAs in flight per warp before their results are scaled (4 costs 8 more registers)
#endif

struct GemvArgs {
    const uint8_t *w;        // packed matrix
    int rows, nbp, n_tiles;  // logical rows, padded blocks per row, tiles
    int tile_align;          // tiles are dealt to CTAs in multiples of this (4 for act*mul so A3 blocks stay whole)
    int qt;                  // 0: Q4_0 matrix, 1: Q4_K matrix
    int rb;                  // finished tiles parked in shared memory between two CTA-wide reductions (<= GEMV_RB_MAX; 2 when k_outer)
    int k_outer;             // 1: loop chunks outermost (long K, <= GEMV_MAXT tiles per CTA); 0: tiles outermost (K <= 2 chunks)
    // input
    const float *x;          // IN_NORM: residual stream, n_in floats
    const float *norm_w;
    float eps;
    int plus_one, n_in;
    const uint8_t *act_in;   // IN_ACT: A3 vector in global memory; IN_F32: fp32 vector of n_in elements (put on the grid here)
    // output
    float *y;                // fp32 output (OUT_ACTMUL: optional copy of h)
    const float *bias_or_res;
    uint8_t *act_out;        // OUT_ACTMUL: A3 of h
    int n_out_act;           // elements of the A3 output vector (for its scale area offset)
    int act;
    float *pmax;             // OUT_LOGITS: per-CTA arg-max partials
    int *pidx;
    int row_off;             // OUT_LOGITS: index of row 0 in the whole vocabulary (a tensor-parallel rank multiplies a row window of the LM head)
    const float *xsq_in;     // IN_NORM: partial sums of x^2 left by the kernel that produced x (fixed order => deterministic)
    int xsq_n;
    float *xsq_out;          // OUT_RES: this CTA's partial sum of squares of the rows it wrote, at [blockIdx.x]
    const uint8_t *pf_ptr;   // weights the NEXT kernel will stream: each CTA pulls the head of its share into L2
    uint64_t pf_total;
    uint32_t l2_window;      // bytes the CTA's L2 prefetch cursor runs ahead of its register loads (0: off)
    uint32_t pf_bytes;       // bytes of the next kernel's weights each CTA pulls into L2 (0: off).  IN_ACT kernels (the O projection, resident and
                             // parked behind attention) issue it BEFORE the wait for their input, the others behind their own stream
    long long *dbg;          // optional [grid][8] %globaltimer stamps of thread 0 (development aid, tools/prof_gemv.py)
};
// The three pointers that differ from layer to layer travel beside the block: in the persistent kernel the block itself is a
// __constant__ object shared by all layers (its fields are instruction operands; taking a per-layer copy would put it in local
// memory, whose loads queue behind the weight stream).  The device code reads these three from GemvPtrs, never from GemvArgs.
struct GemvPtrs {
    const uint8_t *w;
    const float *norm_w;
    const float *bias_or_res;
};
struct GemvArgsTp : GemvArgs {
    // ---- tensor-parallel decode (engine.cu: TpLink; one process per GPU, exchange blocks mapped into each other over NVLink).
    // OUT_TPPART: this rank multiplied its share of K, so its rows are partial sums.  Every row goes, as an 8-byte pair
    // (value, tag of the round), into slot [this rank] of EVERY rank's exchange block (tp_dst[p]: peer memory, one store per
    // rank; an aligned 8-byte store arrives whole).  There is no fence, no counter and no second message: a pair whose tag is
    // the round's tag IS the arrived value.  IN_TPNORM: each warp of the consumer reads the pairs of its own columns from this
    // GPU's block until all carry the tag and forms x = res + slot_0 + ... + slot_{n-1} in that order on every rank, so all
    // ranks hold bit-identical x; CTA 0 also writes x to tp_xout.  The all-reduce is the pair (epilogue of one GEMV, prologue of
    // the next): no kernel of its own, and one NVLink crossing of latency.
    uint2 *tp_dst[8];
    int tp_n;
    const float *tp_res;                 // residual stream before the exchange (n_in floats)
    const uint2 *tp_slots;               // local exchange buffer of this round: slot r at r * tp_slot_stride pairs
    int tp_slot_stride;
    const unsigned long long *tp_fwd;    // forward passes completed so far on this engine (device counter)
    int tp_round, tp_rounds;             // tag of this round = low 32 bits of (*tp_fwd * tp_rounds + tp_round + 1): never 0, never repeated
    float *tp_xout;
    unsigned int *tp_err;                // set to 1 + rank when a peer's pairs did not arrive within TP_TIMEOUT_NS (no hang: the caller reports it)
};
enum { IN_NORM = 0, IN_ACT = 1, IN_F32 = 2, IN_TPNORM = 3 };
enum { OUT_BIAS = 0, OUT_RES = 1, OUT_ACTMUL = 2, OUT_LOGITS = 3, OUT_TPPART = 4 };
#define TP_TIMEOUT_NS 2000000000ll
__device__ __forceinline__ uint4 ld_volatile_v4(const void *p) {   // two (value, tag) pairs, re-read from L2 every time
    uint4 r;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st_volatile_v2(void *p, uint32_t x, uint32_t y) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(x), "r"(y) : "memory");
}
template <class A> __device__ __forceinline__ uint32_t tp_tag(const A &a) {
    return (uint32_t)(__ldcg(a.tp_fwd) * (unsigned long long)a.tp_rounds + (unsigned long long)a.tp_round + 1ull);
}
__device__ __forceinline__ long long global_timer_ns() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// host side (gemv.cu): fill the argument block of one GEMV and report its dynamic shared memory
cudaError_t gemv_make_args(GemvArgs &a, int in_mode, const void *d_packed, int64_t rows, int64_t cols, const float *x, const float *norm_w,
                           float eps, int plus_one, const void *act_in, float *y, const float *bias_or_res, void *act_out, int act,
                           float *pmax, int *pidx, const void *next_w, uint64_t next_bytes, const float *xsq_in, int xsq_n, float *xsq_out,
                           size_t *smem_out, int qt = 0);

// one row of a packed matrix, dequantised (token embedding lookup and tests)
__device__ __forceinline__ float packed_elem(const uint8_t *w, int64_t row, int64_t col, int64_t nbp) {
    PackedLoc loc = packed_loc(row, col / QK, nbp);
    int e = (int)(col % QK);
    uint32_t word = *(const uint32_t *)(w + loc.word0 + ((e & 15) >> 2) * 16);
    int nib = (word >> (8 * (e & 3) + ((e >> 4) << 2))) & 0xF;    // two's complement of q-8
    int v = (nib ^ 8) - 8;
    return (float)v * __half2float(__ushort_as_half(*(const uint16_t *)(w + loc.scale)));
}
// the same for a Q6_K matrix: value = (d * sc) * (q - 32), each product rounded (gguf-py / oracle/quants.py order)
__device__ __forceinline__ float packed_elem_6(const uint8_t *w, int64_t row, int64_t col, int64_t nbp) {
    PackedLoc6 loc = packed_loc_6(row, col / QK, nbp);
    const int e = (int)(col % QK), t = (e & 15) >> 2, kk = e & 3, half = e >> 4;
    const uint32_t word = *(const uint32_t *)(w + loc.word0 + t * 16);
    const int lo = (((word >> (8 * kk + 4 * half)) & 0xF) ^ 8);                 // stored as (q & 15) - 8
    const uint32_t hb = *(const uint16_t *)(w + loc.hi + t * 16);
    const int hi = (hb >> (8 * half + 2 * kk)) & 3;
    const int q = (lo | (hi << 4)) - 32;
    const float d = __half2float(*(const __half *)(w + loc.d));
    const float dsc = __fmul_rn(d, (float)(int)(int8_t)w[loc.sc + half]);
    return __fmul_rn(dsc, (float)q);
}
// the same for a Q4_K matrix: d1 = d*sc, m1 = dmin*m, value = q*d1 - m1, each op rounded (gguf-py / oracle/quants.py order)
__device__ __forceinline__ float packed_elem_k(const uint8_t *w, int64_t row, int64_t col, int64_t nbp) {
    PackedLocK loc = packed_loc_k(row, col / QK, nbp);
    int e = (int)(col % QK);
    uint32_t word = *(const uint32_t *)(w + loc.word0 + ((e & 15) >> 2) * 16);
    int nib = (word >> (8 * (e & 3) + ((e >> 4) << 2))) & 0xF;    // two's complement of q-8
    int q = ((nib ^ 8) - 8) + 8;
    const float2 dm = __half22float2(*(const __half2 *)(w + loc.dm));
    const float d1 = __fmul_rn(dm.x, (float)w[loc.scm]), m1 = __fmul_rn(dm.y, (float)w[loc.scm + 1]);
    return __fsub_rn(__fmul_rn((float)q, d1), m1);
}


__device__ __forceinline__ float act_fn(float g, int act) {
    if (act == B2_ACT_SILU) return g / (1.0f + expf(-g));
    float g3 = 0.044715f * g * g * g;
    return 0.5f * g * (1.0f + tanhf(0.7978845608028654f * (g + g3)));
}
__device__ __forceinline__ void cta_sync() { __syncthreads(); }
__device__ __forceinline__ void l2_prefetch(const void *p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
#ifndef STREAM_EVICT_FIRST
#define STREAM_EVICT_FIRST 0
#endif
__device__ __forceinline__ uint4 ldg_stream4(const void *p) {   // weights: read once, keep out of L1
    uint4 r;
#if STREAM_EVICT_FIRST
    // the line is dead once read: mark it first in line for eviction from L2 so that lines the prefetch cursor brought in survive
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(pol));
#else
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
#endif
    return r;
}
__device__ __forceinline__ uint4 ldcg4(const void *p) {         // activations written by the previous kernel: L2
    uint4 r;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// QT = 0: Q4_0 matrix, QT = 1: Q4_K matrix (layout.cuh).
// One block-group of weights for this lane: rows g / g+8, 4 blocks, + their scales (Q4_0: half2 (d_g, d_g+8) per block;
// Q4_K: bytes (sc_g, m_g, sc_g+8, m_g+8) per block and dm = half2 (d, dmin) of row g, of row g+8).
// Q6_K: wa/wb low nibbles, wh the 2 high bits of the same elements, sd int8 (sc_lo, sc_hi) x (row g, row g+8) per block,
// dd = half2 (d of row g, d of row g+8).
template <int QT> struct WPieceT { uint4 wa, wb, sd; };
template <> struct WPieceT<1> { uint4 wa, wb, sd; uint2 dm; };
template <> struct WPieceT<2> { uint4 wa, wb, sd, wh; uint32_t dd; };
// Activation digits (B fragments of 4 blocks) + block scales (already / 16); Q4_K also needs ts = the sum of the gridded
// activations of block 4*bg + t (this lane's share of the "- dmin*m" term, which multiplies a block SUM, not a dot product).
template <int QT> struct BFragT { uint4 b01, b23; float4 as; };
template <> struct BFragT<1> { uint4 b01, b23; float4 as; float ts; };
template <> struct BFragT<2> { uint4 b01, b23; float4 as; float2 ts2; };   // Q6_K: sums of the two 16-element halves
template <int QT> struct WLay {
    static constexpr int BB = QT == 2 ? BLOCK_BYTES_6 : QT == 1 ? BLOCK_BYTES_K : BLOCK_BYTES;   // bytes per block column of a tile
    static constexpr int SC = QT == 2 ? 160 : QT == 1 ? 192 : 128;                                  // scale bytes per block-group
    static constexpr int SC0 = QT == 2 ? 384 : 256;                                                 // scale area offset = cb * SC0
};
// four 2-bit fields of a byte -> the low 2 bits of four bytes (multiplications by 0x1001 and 65 are shift-adds without overlap)
__device__ __forceinline__ uint32_t spread2(uint32_t b) {
    const uint32_t x = (b * 0x1001u) & 0x000F000Fu;
    return (x * 65u) & 0x03030303u;
}
__device__ __forceinline__ void mma_s8_k16(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
// small unsigned byte k of `word` as a float without I2F: 0x4B0000bb is 8388608 + bb
__device__ __forceinline__ float byte_to_float(uint32_t word, int k) {
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7540u | (uint32_t)k)) - 8388608.0f;
}

// Q6_K: q - 32 = (lo - 8) + 16 u - 24 with lo the low nibble and u = 0..3 the high bits.  Per 16-element half of a block (its
// own int8 scale) two m16n8k16 MMAs: A = 16 (lo - 8) (the nibble trick) and A = u; then
// sum (q-32) X = (D_lo + 256 D_u) / 16 - 24 sum X, the last term through the half-block sums of the activations (ts2).
__device__ __forceinline__ void bg_mma_q6(const WPieceT<2> &r, const BFragT<2> &b, float (&acc)[4], float (&off)[2], int t) {
    const uint32_t ua[4] = {r.wa.x, r.wa.y, r.wa.z, r.wa.w}, ub[4] = {r.wb.x, r.wb.y, r.wb.z, r.wb.w};
    const uint32_t bx[4] = {b.b01.x, b.b01.z, b.b23.x, b.b23.z}, by[4] = {b.b01.y, b.b01.w, b.b23.y, b.b23.w};
    const uint32_t sdw[4] = {r.sd.x, r.sd.y, r.sd.z, r.sd.w};
    const uint32_t hg[4] = {r.wh.x & 0xFFFFu, r.wh.x >> 16, r.wh.y & 0xFFFFu, r.wh.y >> 16};      // row g, blocks 0..3
    const uint32_t hg8[4] = {r.wh.z & 0xFFFFu, r.wh.z >> 16, r.wh.w & 0xFFFFu, r.wh.w >> 16};     // row g+8
    const float asf[4] = {b.as.x, b.as.y, b.as.z, b.as.w};
    const float2 dd = __half22float2(*(const __half2 *)&r.dd);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t sw = sdw[j] ^ 0x80808080u;   // int8 + 128 as unsigned bytes
        const float s0g = __fmul_rn(dd.x, byte_to_float(sw, 0) - 128.0f), s1g = __fmul_rn(dd.x, byte_to_float(sw, 1) - 128.0f);
        const float s0h = __fmul_rn(dd.y, byte_to_float(sw, 2) - 128.0f), s1h = __fmul_rn(dd.y, byte_to_float(sw, 3) - 128.0f);
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            int cl[4] = {MMA_MAGIC, MMA_MAGIC, MMA_MAGIC, MMA_MAGIC}, ch[4] = {MMA_MAGIC, MMA_MAGIC, MMA_MAGIC, MMA_MAGIC};
            const uint32_t al0 = hf ? (ua[j] & 0xF0F0F0F0u) : ((ua[j] << 4) & 0xF0F0F0F0u);
            const uint32_t al1 = hf ? (ub[j] & 0xF0F0F0F0u) : ((ub[j] << 4) & 0xF0F0F0F0u);
            const uint32_t ah0 = spread2((hg[j] >> (8 * hf)) & 0xFFu), ah1 = spread2((hg8[j] >> (8 * hf)) & 0xFFu);
            const uint32_t bb = hf ? by[j] : bx[j];
            mma_s8_k16(cl, al0, al1, bb);
            mma_s8_k16(ch, ah0, ah1, bb);
            const float f0 = (hf ? s1g : s0g) * asf[j], f1 = (hf ? s1h : s0h) * asf[j];
            acc[0] = fmaf(fmaf(256.0f, __int_as_float(ch[0]) - MMA_MAGIC_F, __int_as_float(cl[0]) - MMA_MAGIC_F), f0, acc[0]);
            acc[1] = fmaf(fmaf(256.0f, __int_as_float(ch[1]) - MMA_MAGIC_F, __int_as_float(cl[1]) - MMA_MAGIC_F), f0, acc[1]);
            acc[2] = fmaf(fmaf(256.0f, __int_as_float(ch[2]) - MMA_MAGIC_F, __int_as_float(cl[2]) - MMA_MAGIC_F), f1, acc[2]);
            acc[3] = fmaf(fmaf(256.0f, __int_as_float(ch[3]) - MMA_MAGIC_F, __int_as_float(cl[3]) - MMA_MAGIC_F), f1, acc[3]);
        }
        if (j == t) {   // this lane's block of the "- 24 sum X" term
            off[0] = fmaf(-24.0f, fmaf(s0g, b.ts2.x, s1g * b.ts2.y), off[0]);
            off[1] = fmaf(-24.0f, fmaf(s0h, b.ts2.x, s1h * b.ts2.y), off[1]);
        }
    }
}
// off[0..1] (Q4_K): running sums of (8A - B) * block sum for rows g / g+8, this lane's block (j = t) of every block-group
template <int QT>
__device__ __forceinline__ void bg_mma(const WPieceT<QT> &r, const BFragT<QT> &b, float (&acc)[4], float (&off)[2], int t) {
#ifdef GEMV_EXP_NOMMA
    if (true) { acc[0] += __uint_as_float((r.wa.x ^ r.wb.y ^ r.sd.z) & 0x3fffffu); return; }   // experiment: the loop without its arithmetic
#endif
    if constexpr (QT == 2) {
        bg_mma_q6(r, b, acc, off, t);
    } else {
    const uint32_t ua[4] = {r.wa.x, r.wa.y, r.wa.z, r.wa.w}, ub[4] = {r.wb.x, r.wb.y, r.wb.z, r.wb.w};
    const uint32_t bx[4] = {b.b01.x, b.b01.z, b.b23.x, b.b23.z}, by[4] = {b.b01.y, b.b01.w, b.b23.y, b.b23.w};
    const uint32_t sdw[4] = {r.sd.x, r.sd.y, r.sd.z, r.sd.w};
    const float asf[4] = {b.as.x, b.as.y, b.as.z, b.as.w};
#pragma unroll
    for (int jh = 0; jh < 4; jh += MMA_GROUP) {   // MMA_GROUP independent MMAs at a time, then their scaling into the fp32 accumulators
        int c[MMA_GROUP][4];
#pragma unroll
        for (int u = 0; u < MMA_GROUP; ++u) {
            const int j = jh + u;
            c[u][0] = c[u][1] = c[u][2] = c[u][3] = MMA_MAGIC;
            mma_s8(c[u], (ua[j] << 4) & 0xF0F0F0F0u, (ub[j] << 4) & 0xF0F0F0F0u, ua[j] & 0xF0F0F0F0u, ub[j] & 0xF0F0F0F0u, bx[j], by[j]);
        }
#pragma unroll
        for (int u = 0; u < MMA_GROUP; ++u) {
            const int j = jh + u;
            float2 dd;
            if constexpr (QT == 0) {
                dd = __half22float2(*(const __half2 *)&sdw[j]);
            } else {   // A = d * sc per row (fp32, as the oracle's dequantisation)
                const float2 dg = __half22float2(*(const __half2 *)&r.dm.x), dg8 = __half22float2(*(const __half2 *)&r.dm.y);
                dd.x = __fmul_rn(dg.x, byte_to_float(sdw[j], 0));
                dd.y = __fmul_rn(dg8.x, byte_to_float(sdw[j], 2));
            }
            const float f0 = dd.x * asf[j], f1 = dd.y * asf[j];
            acc[0] = fmaf(__int_as_float(c[u][0]) - MMA_MAGIC_F, f0, acc[0]);
            acc[1] = fmaf(__int_as_float(c[u][1]) - MMA_MAGIC_F, f0, acc[1]);
            acc[2] = fmaf(__int_as_float(c[u][2]) - MMA_MAGIC_F, f1, acc[2]);
            acc[3] = fmaf(__int_as_float(c[u][3]) - MMA_MAGIC_F, f1, acc[3]);
        }
    }
    if constexpr (QT == 1) {   // w = A*(q-8) + (8A - B): the second term times the block's sum of activations, block j = t
        const uint32_t w = t == 0 ? r.sd.x : t == 1 ? r.sd.y : t == 2 ? r.sd.z : r.sd.w;
        const float2 dg = __half22float2(*(const __half2 *)&r.dm.x), dg8 = __half22float2(*(const __half2 *)&r.dm.y);
        const float c0 = __fsub_rn(8.0f * __fmul_rn(dg.x, byte_to_float(w, 0)), __fmul_rn(dg.y, byte_to_float(w, 1)));
        const float c1 = __fsub_rn(8.0f * __fmul_rn(dg8.x, byte_to_float(w, 2)), __fmul_rn(dg8.y, byte_to_float(w, 3)));
        off[0] = fmaf(c0, b.ts, off[0]);
        off[1] = fmaf(c1, b.ts, off[1]);
    }
    }
}
// one piece of the stream: block-group column `warp` of the chunk at `ch` (cb blocks)
__device__ __forceinline__ uint2 ldg_stream2(const void *p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
template <int QT>
__device__ __forceinline__ void load_piece(const uint8_t *ch, int cb, int warp, int lane, WPieceT<QT> &r) {
    r.wa = ldg_stream4(ch + (size_t)warp * 1024 + lane * 16);
    r.wb = ldg_stream4(ch + (size_t)warp * 1024 + 512 + lane * 16);
    r.sd = ldg_stream4(ch + (size_t)cb * WLay<QT>::SC0 + warp * WLay<QT>::SC + (lane >> 2) * 16);
    if constexpr (QT == 1) r.dm = ldg_stream2(ch + (size_t)cb * 256 + warp * WLay<QT>::SC + 128 + (lane >> 2) * 8);
    if constexpr (QT == 2) {
        r.wh = ldg_stream4(ch + (size_t)cb * 256 + (size_t)warp * 512 + lane * 16);
        r.dd = __ldg((const uint32_t *)(ch + (size_t)cb * 384 + warp * 160 + 128 + (lane >> 2) * 4));
    }
}

// shared memory map (bytes): [header 512: 64 floats of scratch, the input-copy mbarrier at +256][red rb*16*16*4*4 = rb*4096]
//                            [A3 of the input nbp*100 (+ fp32 staging)]     rb = GemvArgs::rb, tiles parked between two reductions
#define GEMV_HDR_SMEM 512
#define GEMV_RB_MAX 16                       // one warp finalises one parked tile
#define GEMV_FIXED_SMEM(rb) (GEMV_HDR_SMEM + (rb) * 4096)

// ---- Input staging.  What the previous kernel wrote reaches shared memory through cp.async.bulk (the TMA path): an ordinary load
// would complete only after the ~70 KB of weight loads this SM already has in flight (the L1 load queue is a FIFO).  Thread 0 issues
// the copies after griddepcontrol.wait; every thread waits on the mbarrier, no CTA barrier.  Then the work is warp-local: warp w only
// ever multiplies block-group column w of every chunk, i.e. the input elements [128*kbg, 128*kbg+128) with kbg = ck*16 + w.  It
// normalises exactly those (IN_NORM: the sum of squares arrives as per-CTA partials from the producer of x, added in a fixed order;
// op order of oracle/model.py:_norm), puts them on the activation grid (a 32-element block is 8 consecutive lanes: three shuffles
// for the block maximum) and writes its own block-groups of the shared-memory activation image.
// Shared memory of the input area: [A3 image nbp*100][staging: fp32 vector n_in*4][xsq partials 1 KB]
// Q4_K adds float ts[nbp] after the scales: ts[b] = sum of the gridded activations of block b (staging area moves up by nbp*4).
template <int IN, int QT, class A>
__device__ __forceinline__ void stage_warp_input(const A &a, const GemvPtrs &pv, uint8_t *sact, uint64_t *bar, uint32_t parity, int cpt, int warp, int lane,
                                                 const float4 (&wn)[2], float *scratch, uint32_t tptag) {
    constexpr int A3S = 100 + 4 * QT;   // Q4_K: one block sum per block after the scales, Q6_K: two half-block sums
    float *sts = (float *)(sact + (size_t)a.nbp * 100);
    if (IN == IN_ACT) {   // already on the grid: the image itself was copied
        mbar_wait(bar, parity);
        if constexpr (QT != 0) {   // block sums from the digits: lane = (t, j, half) reads one word of each plane of its block-group
            const float *sscale = (const float *)(sact + (size_t)a.nbp * 96);
            for (int ck = 0; ck < cpt; ++ck) {
                const int kbg = ck * (KC / 4) + warp;
                if (kbg * 4 >= a.nbp) break;
                const uint32_t *pl = (const uint32_t *)(sact + (size_t)kbg * 384) + lane;
                int xs = __dp4a((int)pl[0], 0x01010101, 0) + 128 * __dp4a((int)pl[32], 0x01010101, 0) + 16384 * __dp4a((int)pl[64], 0x01010101, 0);
                if (QT == 1) xs += __shfl_xor_sync(0xffffffffu, xs, 1);   // Q6_K keeps the two 16-element halves apart
                xs += __shfl_xor_sync(0xffffffffu, xs, 8);
                xs += __shfl_xor_sync(0xffffffffu, xs, 16);
                const int j = (lane >> 1) & 3;
                if (QT == 1) { if ((lane & 0x19) == 0) sts[kbg * 4 + j] = (float)xs * sscale[kbg * 4 + j]; }
                else if ((lane & 0x18) == 0) sts[(kbg * 4 + j) * 2 + (lane & 1)] = (float)xs * sscale[kbg * 4 + j];
            }
            __syncwarp();
        }
        return;
    }
    const float *sx = (const float *)(sact + (size_t)a.nbp * A3S);
    const float *sq = sx + a.n_in;
    float *sscale = (float *)(sact + (size_t)a.nbp * 96);
    mbar_wait(bar, parity);
#ifdef GEMV_EXP_NOSTAGE
    if (true) {   // what-if experiment: the input arrives, nothing is normalised or gridded: an all-zero image (wrong results, valid tokens)
        for (int ck = 0; ck < cpt; ++ck) {
            const int kbg = ck * (KC / 4) + warp;
            if (kbg * 4 >= a.nbp) break;
            const int q = kbg * 32 + lane, b = q >> 3, k = q & 7;
            uint8_t *dst = sact + a3_off(b, 0, 4 * k);
            *(uint32_t *)dst = 0u; *(uint32_t *)(dst + 128) = 0u; *(uint32_t *)(dst + 256) = 0u;
            if (k == 0) sscale[b] = 0.f;
        }
        __syncwarp();
        return;
    }
#endif
    float r = 1.0f;
    if constexpr (IN == IN_TPNORM) {
        // x = res + slot_0 + ... : this warp's columns, in place; the sum of squares needs every warp's share.
        // Four ranks' pairs are requested at a time (32 bytes per lane and rank) and re-read until each carries the round's tag.
        float4 *sx4 = (float4 *)(sact + (size_t)a.nbp * A3S);
        const int n4s = a.n_in >> 2;
        const uint32_t tag = tptag;
        const long long t_begin = global_timer_ns();
        bool give_up = false;
        float ss = 0.f;
        for (int ck = 0; ck < cpt; ++ck) {
            const int kbg = ck * (KC / 4) + warp;
            if (kbg * 4 >= a.nbp) break;
            const int q = kbg * 32 + lane;
            if (q < n4s) {
                float4 v = sx4[q];
                for (int p0 = 0; p0 < a.tp_n; p0 += 4) {
                    uint4 lo[4], hi[4];
                    uint32_t spins = 0;
                    for (;;) {
                        bool ok = true;
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            if (p0 + u < a.tp_n) {
                                const uint4 *sp = (const uint4 *)(a.tp_slots + (size_t)(p0 + u) * a.tp_slot_stride) + 2 * q;
                                lo[u] = ld_volatile_v4(sp);
                                hi[u] = ld_volatile_v4(sp + 1);
                            }
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            if (p0 + u < a.tp_n) ok = ok && lo[u].y == tag && lo[u].w == tag && hi[u].y == tag && hi[u].w == tag;
                        if (ok || give_up) break;
                        if ((++spins & 255u) == 0u && (__ldcg(a.tp_err) != 0u || global_timer_ns() - t_begin > TP_TIMEOUT_NS)) {
                            atomicCAS(a.tp_err, 0u, 1u + (unsigned)p0);
                            give_up = true;   // a peer is missing: finish with what is there, the call reports the failure
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (p0 + u < a.tp_n) {
                            v.x = __fadd_rn(v.x, __uint_as_float(lo[u].x)); v.y = __fadd_rn(v.y, __uint_as_float(lo[u].z));
                            v.z = __fadd_rn(v.z, __uint_as_float(hi[u].x)); v.w = __fadd_rn(v.w, __uint_as_float(hi[u].z));
                        }
                }
                sx4[q] = v;
                if (blockIdx.x == 0) __stcg((float4 *)a.tp_xout + q, v);
                ss = fmaf(v.x, v.x, ss); ss = fmaf(v.y, v.y, ss); ss = fmaf(v.z, v.z, ss); ss = fmaf(v.w, v.w, ss);
            }
        }
        ss = warp_sum(ss);
        if (lane == 0) scratch[warp] = ss;
        __syncthreads();
        float tot = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < GEMV_CWARPS; ++w8) tot += scratch[w8];
        r = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(tot, (float)a.n_in), a.eps)));
    }
    if (IN == IN_NORM) {
        float ss = 0.f;
        for (int i = lane; i < a.xsq_n; i += 32) ss += sq[i];
        ss = warp_sum(ss);
        r = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)a.n_in), a.eps)));
    }
    const int n4 = a.n_in >> 2;
    for (int ck = 0; ck < cpt; ++ck) {
        const int kbg = ck * (KC / 4) + warp;
        if (kbg * 4 >= a.nbp) break;
        const int q = kbg * 32 + lane;   // float4 index: block = q / 8, word-in-block = q % 8
        float yv[4] = {0.f, 0.f, 0.f, 0.f};
        if (q < n4) {
            const float4 v = ((const float4 *)sx)[q];
            if (IN == IN_NORM || IN == IN_TPNORM) {
                const float4 w = ck == 0 ? wn[0] : ck == 1 ? wn[1] : __ldg((const float4 *)pv.norm_w + q);   // constant indices: wn stays in registers
                const float xs[4] = {v.x, v.y, v.z, v.w}, ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) yv[e] = __fmul_rn(__fmul_rn(xs[e], r), a.plus_one ? __fadd_rn(1.0f, ws[e]) : ws[e]);
            } else {
                yv[0] = v.x; yv[1] = v.y; yv[2] = v.z; yv[3] = v.w;
            }
        }
        float m = fmaxf(fmaxf(fabsf(yv[0]), fabsf(yv[1])), fmaxf(fabsf(yv[2]), fabsf(yv[3])));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
        const int ex = max((int)((__float_as_uint(m) >> 23) & 0xFF) - 127, -100);
        const float sc = __uint_as_float((uint32_t)(ex - (ACT_BITS - 1) + 127) << 23);
        const float inv = __uint_as_float((uint32_t)((ACT_BITS - 1) - ex + 127) << 23);
        // Balanced base-128 digits of four elements at once (same digits as act_digits, fewer instructions: every CTA grids the
        // whole vector, sixteen warps on four issue ports).  With U = X + 64*(1 + 128 + 16384) > 0 the digits biased by 64 are
        // U & 127, (U >> 7) & 127 and U >> 14; the low bytes of the four U (U >> 7, U >> 14) are gathered with PRMT, and
        // "- 64" on 7-bit fields is a flip of bit 6 followed by its sign extension into bit 7.  The top digit can reach +64
        // (|X| = 2^20 after rounding), so it is not masked but de-biased per element.
        int xsum = 0;
        uint32_t U[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int X = __float2int_rn(yv[e] * inv);
            xsum += X;
            U[e] = (uint32_t)(X + 64 * (1 + 128 + 16384));
        }
        auto gather = [](uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3) {   // byte 0 of each word
            return __byte_perm(__byte_perm(a0, a1, 0x0040), __byte_perm(a2, a3, 0x0040), 0x5410);
        };
        auto debias7 = [](uint32_t w) {   // four 7-bit fields u -> int8 (u - 64)
            const uint32_t tt = (w & 0x7F7F7F7Fu) ^ 0x40404040u;
            return tt | ((tt & 0x40404040u) << 1);
        };
        const uint32_t w0 = debias7(gather(U[0], U[1], U[2], U[3]));
        const uint32_t w1 = debias7(gather(U[0] >> 7, U[1] >> 7, U[2] >> 7, U[3] >> 7));
        const uint32_t w2 = gather((U[0] >> 14) - 64u, (U[1] >> 14) - 64u, (U[2] >> 14) - 64u, (U[3] >> 14) - 64u);
        const int b = q >> 3, k = q & 7;                    // elements 4k..4k+3 of block b
        uint8_t *dst = sact + a3_off(b, 0, 4 * k);
        *(uint32_t *)dst = w0;
        *(uint32_t *)(dst + 128) = w1;
        *(uint32_t *)(dst + 256) = w2;
        if (k == 0) sscale[b] = (q < n4) ? sc : 0.f;
        if constexpr (QT != 0) {
            xsum += __shfl_xor_sync(0xffffffffu, xsum, 1);
            xsum += __shfl_xor_sync(0xffffffffu, xsum, 2);
            if (QT == 1) {
                xsum += __shfl_xor_sync(0xffffffffu, xsum, 4);
                if (k == 0) sts[b] = (q < n4) ? (float)xsum * sc : 0.f;
            } else if ((k & 3) == 0) {   // lanes k = 0..3 hold elements 0..15, k = 4..7 elements 16..31
                sts[2 * b + (k >> 2)] = (q < n4) ? (float)xsum * sc : 0.f;
            }
        }
    }
    __syncwarp();
}

// ---- reduction + epilogue of the `batch_n` finished tiles parked in shared memory (called by all threads).
// Every warp parked its K-partial of the 16 rows x 3 digit planes; warp w finalises tile batch_first + w.
// Two CTA-wide barriers per a.rb tiles instead of one per tile: the weight streams of the sixteen warps are not
// re-synchronised on memory latency every other piece.
// What the epilogue needs from the argument block, by value: a reference to the kernel parameter would force the whole block
// into local memory of every thread (this function is not inlined).
template <bool TP> struct FlushArgs { float *y; int rows, act, row_off; };
template <> struct FlushArgs<true> { float *y; int rows, act, row_off; uint2 *tp_dst[8]; int tp_n; uint32_t tp_tag; };
template <int OUT, class A>
__device__ __forceinline__ FlushArgs<OUT == OUT_TPPART> make_flush_args(const A &a, uint32_t tptag) {
    FlushArgs<OUT == OUT_TPPART> f;
    f.y = a.y; f.rows = a.rows; f.act = a.act; f.row_off = a.row_off;
    if constexpr (OUT == OUT_TPPART) {
#pragma unroll
        for (int p = 0; p < 8; ++p) f.tp_dst[p] = a.tp_dst[p];
        f.tp_n = a.tp_n;
        f.tp_tag = tptag;
    }
    return f;
}
struct FlushRun { float sumsq; };   // running sum of squares of the rows this thread finalised (OUT_RES)
template <int OUT, int QT>
__device__ __noinline__ FlushRun flush_batch(const FlushArgs<OUT == OUT_TPPART> a, const float *bias_or_res, float *red, int batch_first, int batch_n, FlushRun run) {
    float sumsq = run.sumsq;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    __syncthreads();
    if (warp < batch_n) {
        // lane = (half, row): each lane adds 8 of the 16 warp partials of its row in a fixed order, the halves meet by shuffle
        const int tile = batch_first + warp;
        const float *rb = red + (size_t)warp * (GEMV_CWARPS * 16 * 4);
        const int r = lane & 15, hf = lane >> 4;
        float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < GEMV_CWARPS / 2; ++w8) {
            const float4 p4 = *(const float4 *)(rb + ((hf * (GEMV_CWARPS / 2) + w8) * 16 + r) * 4);
            v0 += p4.x;
            v1 += p4.y;
            v2 += p4.z;
            if (QT != 0) v3 += p4.w;   // Q4_K / Q6_K: the term that multiplies block sums of the activations
        }
        v0 += __shfl_xor_sync(0xffffffffu, v0, 16);
        v1 += __shfl_xor_sync(0xffffffffu, v1, 16);
        v2 += __shfl_xor_sync(0xffffffffu, v2, 16);
        float yv = fmaf(16384.0f, v2, fmaf(128.0f, v1, v0));
        if (QT != 0) yv += v3 + __shfl_xor_sync(0xffffffffu, v3, 16);
        const int row = tile * TILE_ROWS + r;
        if (OUT == OUT_BIAS) {
            if (lane < 16 && row < a.rows) a.y[row] = bias_or_res ? yv + bias_or_res[row] : yv;
        } else if (OUT == OUT_RES) {
            if (lane < 16 && row < a.rows) {
                const float xn = __ldcg(bias_or_res + row) + yv;
                a.y[row] = xn;
                sumsq = fmaf(xn, xn, sumsq);
            }
        } else if constexpr (OUT == OUT_TPPART) {
            if (lane < 16 && row < a.rows) {   // this rank's partial sum into its slot on every rank (16 lanes = 128 contiguous bytes per rank)
#pragma unroll
                for (int p = 0; p < 8; ++p)
                    if (p < a.tp_n) st_volatile_v2(a.tp_dst[p] + row, __float_as_uint(yv), a.tp_tag);
            }
        } else if (OUT == OUT_LOGITS) {
            // the running arg-max lives in shared memory, one (value, index) per warp in the header in front of the parked tiles (scratch[32..47],
            // scratch[48..63]; the first 32 floats serve the staging): two registers less across the whole stream loop (this variant spilled four around every block-group)
            float bv = -INFINITY;
            int bi = 0x7fffffff;
            if (lane < 16 && row < a.rows) {
                a.y[row] = yv;
                bv = yv;
                bi = row + a.row_off;
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (lane == 0) {
                float *sv = (float *)((uint8_t *)red - GEMV_HDR_SMEM) + 32;
                int *si = (int *)sv + 16;
                if (bv > sv[warp] || (bv == sv[warp] && bi < si[warp])) { sv[warp] = bv; si[warp] = bi; }
            }
        } else {   // OUT_ACTMUL: rows (2j, 2j+1) = (gate_j, up_j)
            const float up = __shfl_down_sync(0xffffffffu, yv, 1);
            const float hv = act_fn(yv, a.act) * up;
            const int pair = tile * 8 + (r >> 1);
            if (lane < 16 && !(r & 1) && 2 * pair < a.rows) a.y[pair] = hv;
        }
    }
    __syncthreads();   // the parked partials may be overwritten from here on
    return FlushRun{sumsq};
}

// One GEMV executed by a whole 512-thread CTA: the body of the kernel k_gemv (ordered after the previous kernel by
// griddepcontrol.wait).
// PATH: 0 / 1 / 2 = tiles-outermost with two chunks per tile / with one chunk / chunks-outermost (-1 = all three compiled in and
//       chosen at run time).  The kernels are instantiated per path: a third of the code per kernel means fewer instruction-cache
//       misses at every kernel start (no_instructions was 4-13 % of the warp samples of these 10-30 us kernels).
template <int IN, int OUT, int PATH = -1, int QT = 0, class A = GemvArgs>
__device__ __forceinline__ void gemv_phase(const A &a, const GemvPtrs pv, uint8_t *smem, uint32_t &bar_parity) {
    float *scratch = (float *)smem;                                             // 64 floats
    uint64_t *bar = (uint64_t *)(smem + 256);                                   // completion of the input copies
    float *red = (float *)(smem + GEMV_HDR_SMEM);                               // [rb tiles][16 warps][16 rows][4]
    uint8_t *sact = smem + GEMV_FIXED_SMEM(a.rb);                               // A3 image of the input vector (+ fp32 staging, partials)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int bid = blockIdx.x;
    const int g = lane >> 2, t = lane & 3;
    if (tid == 0) {
        pdl_launch_dependents();
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    auto stamp = [&](int k) {
        if (a.dbg && tid == 0) {
            long long tt;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
            a.dbg[bid * 8 + k] = tt;
        }
    };
    stamp(0);
    // tensor-parallel modes: the tag of this exchange round (a device counter of forward passes: the graph is replayed unchanged);
    // requested here, first used after the wait / in the epilogue
    uint32_t tptag = 0;
    if constexpr (IN == IN_TPNORM || OUT == OUT_TPPART) tptag = tp_tag(a);

    // contiguous, aligned share of the tiles
    const int units = (a.n_tiles + a.tile_align - 1) / a.tile_align;
    const int u0 = (int)(((uint32_t)units * bid) / gridDim.x), u1 = (int)(((uint32_t)units * (bid + 1)) / gridDim.x);
    const int t0 = u0 * a.tile_align, t1 = min(a.n_tiles, u1 * a.tile_align);
    const int nt = t1 - t0;
    const int cpt = (a.nbp + KC - 1) / KC;                                      // chunks per tile
    const uint8_t *base = pv.w + (uint64_t)t0 * a.nbp * WLay<QT>::BB;
    const uint32_t tile_bytes = (uint32_t)a.nbp * WLay<QT>::BB;
    const int cb_last = a.nbp - (cpt - 1) * KC;                                 // blocks in the last chunk of a tile
    const bool act_full = warp < (KC >> 2);                                     // this warp has a block-group in a full chunk ...
    const bool act_last = warp < (cb_last >> 2);                                // ... in the last chunk of a tile

    // B fragments of block-group column `warp` of chunk ck, from the shared-memory activation image
    auto load_b = [&](int ck, BFragT<QT> &b) {
        const int kbg = ck * (KC / 4) + warp;
        b.b01 = make_uint4(0u, 0u, 0u, 0u);
        b.b23 = make_uint4(0u, 0u, 0u, 0u);
        b.as = make_float4(0.f, 0.f, 0.f, 0.f);
        if (kbg * 4 < a.nbp) {
            if (g < 3) {
                const uint4 *bp = (const uint4 *)(sact + (size_t)kbg * 384 + g * 128 + t * 32);
                b.b01 = bp[0];
                b.b23 = bp[1];
            }
            b.as = *(const float4 *)(sact + (size_t)a.nbp * 96 + kbg * 16);
            b.as.x *= 0.0625f; b.as.y *= 0.0625f; b.as.z *= 0.0625f; b.as.w *= 0.0625f;   // the 1/16 of the nibble trick
        }
        if constexpr (QT == 1) b.ts = (kbg * 4 < a.nbp) ? *(const float *)(sact + (size_t)a.nbp * 100 + (kbg * 4 + t) * 4) : 0.f;
        if constexpr (QT == 2) b.ts2 = (kbg * 4 < a.nbp) ? *(const float2 *)(sact + (size_t)a.nbp * 100 + (kbg * 4 + t) * 8) : make_float2(0.f, 0.f);
    };
    // pull this CTA's share of the next kernel's weights into L2 (lane 0 of every warp, 16 KB per instruction)
    auto prefetch_next = [&]() {
        if (lane == 0 && a.pf_ptr) {
            const uint64_t share = (a.pf_total / gridDim.x) & ~(uint64_t)127;
            const uint8_t *p0 = a.pf_ptr + share * bid;
            const uint64_t n = share < a.pf_bytes ? share : a.pf_bytes;
            for (uint64_t o = (uint64_t)warp * 16384; o + 16384 <= n; o += GEMV_CWARPS * 16384) l2_prefetch(p0 + o, 16384);
        }
    };

    float sumsq = 0.f;
    // The norm weights of this warp's columns are immutable: they are fetched before waiting for the producer of the input.
    float4 wn[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
    // Wait for the producer of the input vector and start its copy into shared memory (thread 0, cp.async.bulk onto `bar`).
    auto input_wait_copy = [&]() {
        stamp(1);
        if (IN == IN_NORM || IN == IN_TPNORM) {
#pragma unroll
            for (int ck = 0; ck < 2; ++ck) {
                const int q = (ck * (KC / 4) + warp) * 32 + lane;
                if (ck < cpt && q < (a.n_in >> 2)) wn[ck] = __ldg((const float4 *)pv.norm_w + q);
            }
        }
        __syncthreads();            // everybody has issued its first weight loads; the mbarrier thread 0 initialised is visible
        if (tid == 0) {
            pdl_wait();             // the copies below read what the previous kernel wrote
            auto rot_copy = [&](uint8_t *dst, const uint8_t *src, uint32_t bytes) { bulk_g2s(dst, src, bytes, bar); };
            if constexpr (IN == IN_TPNORM) {
                // only the residual travels through shared memory; the warps read the exchange slots themselves
                float *sx = (float *)(sact + (size_t)a.nbp * (100 + 4 * QT));
                const uint32_t bytes = (uint32_t)a.n_in * 4;
                mbar_expect_tx(bar, bytes);
                bulk_g2s(sx, a.tp_res, bytes, bar);
            } else if (IN == IN_ACT) {
                const uint32_t bytes = (uint32_t)a.nbp * 100;
                mbar_expect_tx(bar, bytes);
                rot_copy(sact, a.act_in, bytes);
            } else {
                float *sx = (float *)(sact + (size_t)a.nbp * (100 + 4 * QT));
                const uint32_t bytes = (uint32_t)a.n_in * 4;
                const uint32_t qbytes = IN == IN_NORM ? (uint32_t)((a.xsq_n + 3) & ~3) * 4 : 0;
                mbar_expect_tx(bar, bytes + qbytes);
                rot_copy((uint8_t *)sx, IN == IN_NORM ? (const uint8_t *)a.x : (const uint8_t *)a.act_in, bytes);
                if (qbytes) bulk_g2s(sx + a.n_in, a.xsq_in, qbytes, bar);
            }
        }
        stamp(2);   // everything from here reads what the previous kernel wrote (ld.global.cg: under programmatic
                      // dependent launch a CTA of that kernel may have shared this SM's L1)
    };
    // Wait for the copy and put this warp's columns of the input on the activation grid.
    auto input_stage = [&]() {
        stage_warp_input<IN, QT>(a, pv, sact, bar, bar_parity, cpt, warp, lane, wn, scratch, tptag);
        bar_parity ^= 1;
        stamp(3);
    };
    // Order of the first weight requests, the wait and the staging: the requests go out first (weights never depend on the
    // previous kernel) and are in flight while griddepcontrol.wait blocks.  (Shared-memory loads and stores of the staging complete
    // in order behind the global loads this SM has in flight; the other orders measured slower)
    auto stage_input = [&](auto &&first_loads) {
        if (GEMV_ORDER == 1) {   // experiment: staging before the first register loads (they then hit the L2 window)
            input_wait_copy();
            input_stage();
            first_loads();
        } else {
            first_loads();
            if constexpr (IN == IN_ACT) prefetch_next();   // one call site per kernel: a second copy of it costs the main kernels registers (spills)
            input_wait_copy();
            input_stage();
        }
    };
    if (OUT == OUT_LOGITS && lane == 0) { scratch[32 + warp] = -INFINITY; ((int *)scratch)[48 + warp] = 0x7fffffff; }   // read and written by the same lane only, in flush_batch
    int batch_first = t0, batch_n = 0;
    // only the residual epilogue carries a value across flushes in registers (its sum of squares); the arg-max of the LM head lives in
    // shared memory.  Nothing else travels through the call: what does costs registers through the whole stream loop.
    auto do_flush = [&]() {
        if constexpr (OUT == OUT_RES) sumsq = flush_batch<OUT, QT>(make_flush_args<OUT>(a, tptag), pv.bias_or_res, red, batch_first, batch_n, FlushRun{sumsq}).sumsq;
        else (void)flush_batch<OUT, QT>(make_flush_args<OUT>(a, tptag), pv.bias_or_res, red, batch_first, batch_n, FlushRun{0.f});
    };
    auto park = [&](float (&acc)[4], float (&off)[2]) {   // park the finished tile's partial; reduce when a.rb tiles are waiting
        float *rb = red + (size_t)batch_n * (GEMV_CWARPS * 16 * 4);
        if constexpr (QT != 0) {   // the four t lanes covered one block each: their sum rides in the unused fourth column (plane "3")
            float o0 = off[0], o1 = off[1];
            o0 += __shfl_xor_sync(0xffffffffu, o0, 1); o0 += __shfl_xor_sync(0xffffffffu, o0, 2);
            o1 += __shfl_xor_sync(0xffffffffu, o1, 1); o1 += __shfl_xor_sync(0xffffffffu, o1, 2);
            if (t == 1) { acc[1] = o0; acc[3] = o1; }
            off[0] = off[1] = 0.f;
        }
        if (t < 2) {
            *(float2 *)(rb + (warp * 16 + g) * 4 + 2 * t) = make_float2(acc[0], acc[1]);
            *(float2 *)(rb + (warp * 16 + g + 8) * 4 + 2 * t) = make_float2(acc[2], acc[3]);
        }
        acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
        if (++batch_n == a.rb) {
            do_flush();
            batch_first += batch_n;
            batch_n = 0;
        }
    };
    // GEMV_DEPTH = 4 pieces in flight per warp.  Defined on every path (a warp without work in a piece skips the load): a register
    // that is read-before-written on some path counts as live around the enclosing token/layer loops of the persistent kernel,
    // i.e. through every OTHER stage, and the stages then spill each other's pipelines.
    const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
    WPieceT<QT> buf0, buf1, buf2, buf3;
    buf0.wa = buf0.wb = buf0.sd = z4; buf1 = buf0; buf2 = buf0; buf3 = buf0;
    if constexpr (QT == 1) { buf0.dm = make_uint2(0u, 0u); buf1.dm = buf0.dm; buf2.dm = buf0.dm; buf3.dm = buf0.dm; }
    if constexpr (QT == 2) { buf0.wh = z4; buf0.dd = 0u; buf1 = buf0; buf2 = buf0; buf3 = buf0; }
    float off[2] = {0.f, 0.f}, off1[2] = {0.f, 0.f};   // Q4_K: the "- dmin*m" sums of the tile(s) in flight
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    if (PATH == -1 ? !a.k_outer : PATH != 2) {
        // ================= tiles outermost: the CTA's bytes are one contiguous run of chunks; K spans one or two chunks whose
        // B fragments stay in registers for the whole kernel.  With two chunks per tile the even pieces are full chunks and
        // the odd ones the (possibly shorter) last chunk of a tile: no index arithmetic in the loop.
        const uint8_t *lp = base;
        int l_rem = nt * cpt;            // pieces not requested yet
        // L2 prefetch cursor (thread 0): cp.async.bulk.prefetch.L2 keeps the CTA's contiguous byte range L2_WINDOW_BYTES ahead of
        // the register loads, so those hit L2 (short latency) while HBM streams at its own pace from the first cycle
        // (lane 0 of warp w owns the 16 KB blocks w, w+16, ... of the range: one instruction per warp per 256 KB of progress)
        const uint8_t *pf = base + (size_t)warp * 16384, *pf_end = base + (size_t)nt * tile_bytes;
        auto l2_window = [&]() {
            if (lane == 0 && pf < pf_end && pf < lp + a.l2_window) {
                const size_t left = (size_t)(pf_end - pf);
                l2_prefetch(pf, left < 16384 ? (uint32_t)left & ~15u : 16384u);
                pf += GEMV_CWARPS * 16384;
            }
        };
        if (GEMV_ORDER != 2) l2_window();
        auto load_full = [&](WPieceT<QT> &r) {
            if (l_rem > 0) { if (act_full) load_piece<QT>(lp, KC, warp, lane, r); lp += KC * WLay<QT>::BB; --l_rem; }
        };
        auto load_last = [&](WPieceT<QT> &r) {
            if (l_rem > 0) { if (act_last) load_piece<QT>(lp, cb_last, warp, lane, r); lp += cb_last * WLay<QT>::BB; --l_rem; }
        };
        if (PATH == -1 ? cpt == 2 : PATH == 0) {
            stage_input([&]() { load_full(buf0); load_last(buf1); load_full(buf2); load_last(buf3); });
            BFragT<QT> bf0, bf1;
            load_b(0, bf0);
            load_b(1, bf1);
            for (int rem = nt; rem > 0; rem -= 2) {
                if (act_full) bg_mma<QT>(buf0, bf0, acc, off, t);
                load_full(buf0);
                if (act_last) bg_mma<QT>(buf1, bf1, acc, off, t);
                load_last(buf1);
                park(acc, off);
                if (rem > 1) {
                    if (act_full) bg_mma<QT>(buf2, bf0, acc, off, t);
                    load_full(buf2);
                    if (act_last) bg_mma<QT>(buf3, bf1, acc, off, t);
                    load_last(buf3);
                    park(acc, off);
                }
                l2_window();
            }
        } else {
            stage_input([&]() { load_last(buf0); load_last(buf1); load_last(buf2); load_last(buf3); });
            BFragT<QT> bf0;
            load_b(0, bf0);
            for (int rem = nt; rem > 0; rem -= 4) {
                if (act_last) bg_mma<QT>(buf0, bf0, acc, off, t);
                load_last(buf0);
                park(acc, off);
                if (rem > 1) { if (act_last) bg_mma<QT>(buf1, bf0, acc, off, t); load_last(buf1); park(acc, off); }
                if (rem > 2) { if (act_last) bg_mma<QT>(buf2, bf0, acc, off, t); load_last(buf2); park(acc, off); }
                if (rem > 3) { if (act_last) bg_mma<QT>(buf3, bf0, acc, off, t); load_last(buf3); park(acc, off); }
                l2_window();
            }
        }
    } else {
        // ================= chunks outermost (K longer than two chunks; a CTA owns one or two tiles whose accumulators stay in
        // registers).  A step is chunk ck of both tiles; the B fragments of a step are read from shared memory one step ahead.
        float acc1[4] = {0.f, 0.f, 0.f, 0.f};
        const bool two = nt > 1;
        int l_ck = 0;                     // chunk of the next step to request
        auto load_step = [&](WPieceT<QT> &r0, WPieceT<QT> &r1) {
            if (l_ck < cpt) {
                const bool last = l_ck == cpt - 1;
                if (last ? act_last : act_full) {
                    const uint8_t *ch = base + (size_t)l_ck * (KC * WLay<QT>::BB);
                    load_piece<QT>(ch, last ? cb_last : KC, warp, lane, r0);
                    if (two) load_piece<QT>(ch + tile_bytes, last ? cb_last : KC, warp, lane, r1);
                }
                ++l_ck;
            }
        };
        auto l2_all = [&]() {
            if (lane == 0 && a.l2_window) {   // a CTA's one or two tiles (<= 2 x 220 KB) go to L2 at once, a 16 KB block per warp and round
                for (uint32_t o = (uint32_t)warp * 16384; o + 16384 <= (uint32_t)nt * tile_bytes; o += GEMV_CWARPS * 16384) l2_prefetch(base + o, 16384);
            }
        };
        if (GEMV_ORDER != 2) l2_all();
        stage_input([&]() { load_step(buf0, buf1); load_step(buf2, buf3); });
        if (GEMV_ORDER == 2) l2_all();
        BFragT<QT> bcur, bnxt;
        load_b(0, bcur);
        for (int ck = 0; ck < cpt; ck += 2) {
            load_b(ck + 1, bnxt);
            if ((ck == cpt - 1) ? act_last : act_full) {
                bg_mma<QT>(buf0, bcur, acc, off, t);
                if (two) bg_mma<QT>(buf1, bcur, acc1, off1, t);
            }
            load_step(buf0, buf1);
            if (ck + 1 < cpt) {
                load_b(ck + 2, bcur);
                if ((ck + 1 == cpt - 1) ? act_last : act_full) {
                    bg_mma<QT>(buf2, bnxt, acc, off, t);
                    if (two) bg_mma<QT>(buf3, bnxt, acc1, off1, t);
                }
                load_step(buf2, buf3);
            }
        }
        if (nt > 0) park(acc, off);
        if (two) park(acc1, off1);
    }
    stamp(4);
    // This CTA has requested all of its own weights: the head of its share of the NEXT kernel's weights starts flowing HBM -> L2
    // now, behind the stream, and fills the time HBM would idle through this kernel's tail and the hand-over to the next one.
    if constexpr (IN != IN_ACT) prefetch_next();
    if (batch_n > 0) {
        do_flush();
    }
    stamp(5);
    if (OUT == OUT_RES && a.xsq_out) {   // partial sum of squares of the rows this CTA wrote (input of the next RMSNorm)
        sumsq = warp_sum(sumsq);
        __syncthreads();
        if (lane == 0) scratch[warp] = sumsq;
        __syncthreads();
        if (tid == 0) {
            float tot = 0.f;
#pragma unroll
            for (int w8 = 0; w8 < GEMV_CWARPS; ++w8) tot += scratch[w8];
            a.xsq_out[bid] = tot;
        }
    }
    if (OUT == OUT_LOGITS) {   // every warp finalised some tiles: merge the sixteen running maxima flush_batch left in shared memory
        float *sv = scratch + 32;
        int *si = (int *)(scratch + 48);
        __syncthreads();
        if (warp == 0) {
            float best_v = lane < GEMV_CWARPS ? sv[lane] : -INFINITY;
            int best_i = lane < GEMV_CWARPS ? si[lane] : 0x7fffffff;
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                float ov = __shfl_xor_sync(0xffffffffu, best_v, o);
                int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
                if (ov > best_v || (ov == best_v && oi < best_i)) { best_v = ov; best_i = oi; }
            }
            if (lane == 0) { a.pmax[bid] = best_v; a.pidx[bid] = best_i; }
        }
    }
}

