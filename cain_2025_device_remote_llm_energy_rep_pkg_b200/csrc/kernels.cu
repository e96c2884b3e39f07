// kernels.cu — hand-written sm_100a kernels of the batch-1 decode path.
// Stage map (Ollama call site experiment/RunnerConfig.py:128-131):
//   embed_q4_0        a3.1  token embedding row (dequant, gemma sqrt(H) scale)
//   rmsnorm_quant     a3.2  RMSNorm (gemma 1+w) fused with the activation grid
//   gemv_q4_0         a3.3/a3.6/a3.7/a3.8  QKV(+bias) / O(+res) / gate,up -> act*mul / down(+res) / LM head
//   rope_kv           a3.4  RoPE + f16 append into the paged KV cache
//   attn_decode       a3.5  paged attention over the cache (GQA/MQA/MHA), split over the context
//   argmax            a3.8  greedy sampler
#include "kernels.cuh"
#include "layout.cuh"

// =============================================================================== activation grid
__global__ void k_quant_act(const float *__restrict__ x, int64_t n, void *__restrict__ act) {
    int lane = threadIdx.x & 31;
    int64_t nb = n / QK;
    for (int64_t b = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32; b < nb; b += (int64_t)gridDim.x * (blockDim.x / 32)) {
        float s;
        int X = act_grid_lane(x[b * QK + lane], s);
        a3_store_block(act, n, b, lane, X, s);
    }
}
void launch_quant_act(const float *d_x, int64_t n, void *d_act, cudaStream_t st) {
    int64_t nb = n / QK;
    unsigned grid = (unsigned)((nb + 7) / 8);
    k_quant_act<<<grid, 256, 0, st>>>(d_x, n, d_act);
}

// RMSNorm fused with the activation grid; one CTA (n <= 16384).  Op order follows oracle/model.py:_norm:
//   r = 1/sqrt(mean(x^2) + eps);  y = (x * r) * w   or   (x * r) * (1 + w)
__global__ void __launch_bounds__(1024) k_rmsnorm_quant(const float *__restrict__ x, const float *__restrict__ w, int n, float eps,
                                                        int plus_one, void *__restrict__ act, float *__restrict__ y_opt) {
    __shared__ float red[32];
    __shared__ float s_r;
    float ss = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { float v = x[i]; ss = fmaf(v, v, ss); }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x < 32) {
        float t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
        t = warp_sum(t);
        if (threadIdx.x == 0) s_r = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(t, (float)n), eps)));
    }
    __syncthreads();
    float r = s_r;
    int lane = threadIdx.x & 31;
    for (int b = threadIdx.x >> 5; b < n / QK; b += blockDim.x >> 5) {
        int i = b * QK + lane;
        float wi = w[i];
        if (plus_one) wi = __fadd_rn(1.0f, wi);
        float yv = __fmul_rn(__fmul_rn(x[i], r), wi);
        if (y_opt) y_opt[i] = yv;
        float s;
        int X = act_grid_lane(yv, s);
        a3_store_block(act, n, b, lane, X, s);
    }
}
void launch_rmsnorm_quant(const float *d_x, const float *d_w, int64_t n, float eps, int plus_one, void *d_act, float *d_y_opt, cudaStream_t st) {
    int threads = n >= 4096 ? 1024 : (n >= 1024 ? 512 : 256);
    k_rmsnorm_quant<<<1, threads, 0, st>>>(d_x, d_w, (int)n, eps, plus_one, d_act, d_y_opt);
}

// =============================================================================== RoPE + KV append
// Block h < nh rotates query head h in place; nh <= h < nh+nkv rotates key head and appends it (f16);
// the last nkv blocks append the value head.  Rotate-half pairing (j, j+hd/2); products and sums are
// individually rounded like the oracle's tensor ops (oracle/model.py:_rope).
__global__ void k_rope_kv(float *__restrict__ qkv, int nh, int nkv, int hd, const int *__restrict__ d_pos, int pos_imm,
                          const float *__restrict__ cs, const float *__restrict__ sn, __half *__restrict__ kc, __half *__restrict__ vc,
                          const int32_t *__restrict__ page_table) {
    const int pos = d_pos ? *d_pos : pos_imm;
    const int h = blockIdx.x, half = hd >> 1;
    const int page = page_table[pos / PAGE_TOKENS], slot = pos % PAGE_TOKENS;
    if (h < nh + nkv) {
        float *t = qkv + (int64_t)h * hd;   // q heads then k heads are contiguous in the fused QKV output
        for (int j = threadIdx.x; j < half; j += blockDim.x) {
            float c = cs[(int64_t)pos * half + j], s = sn[(int64_t)pos * half + j];
            float a = t[j], b = t[j + half];
            float ra = __fsub_rn(__fmul_rn(a, c), __fmul_rn(b, s));
            float rb = __fadd_rn(__fmul_rn(b, c), __fmul_rn(a, s));
            if (h < nh) {
                t[j] = ra;
                t[j + half] = rb;
            } else {
                int kvh = h - nh;
                __half *dst = kc + (((int64_t)page * nkv + kvh) * PAGE_TOKENS + slot) * hd;
                dst[j] = __float2half_rn(ra);
                dst[j + half] = __float2half_rn(rb);
            }
        }
    } else {
        int kvh = h - nh - nkv;
        const float *t = qkv + (int64_t)(nh + nkv + kvh) * hd;
        __half *dst = vc + (((int64_t)page * nkv + kvh) * PAGE_TOKENS + slot) * hd;
        for (int j = threadIdx.x; j < hd; j += blockDim.x) dst[j] = __float2half_rn(t[j]);
    }
}
void launch_rope_kv(float *d_qkv, int nh, int nkv, int hd, const int *d_pos, int pos_imm, const float *d_cos, const float *d_sin,
                    __half *d_k, __half *d_v, const int32_t *d_page_table, cudaStream_t st) {
    k_rope_kv<<<nh + 2 * nkv, 128, 0, st>>>(d_qkv, nh, nkv, hd, d_pos, pos_imm, d_cos, d_sin, d_k, d_v, d_page_table);
}

// =============================================================================== paged attention, decode
// grid (nh, ATTN_NSPLIT), 128 threads.  Split s covers pages [s*pps, (s+1)*pps) of the context.
// Phase 1: 8 lanes per key, 16 keys (one page) per iteration, 16-byte f16 loads, scores to smem.
// Phase 2: thread d accumulates sum_t p_t * V[t][d] (and d+128 for hd 256).
// Partials (m, l, o[hd]) go to scratch; the last split to finish for a head merges them (ticket counter).
uint64_t attn_scratch_bytes(int nh, int hd) { return (uint64_t)nh * ATTN_NSPLIT * (hd + 2) * 4 + (uint64_t)nh * 4; }

__global__ void __launch_bounds__(128) k_attn_decode(const float *__restrict__ q, int nh, int nkv, int hd, const int *__restrict__ d_pos,
                                                     int nctx_imm, const __half *__restrict__ kc, const __half *__restrict__ vc,
                                                     const int32_t *__restrict__ page_table, float *__restrict__ out,
                                                     float *__restrict__ part, int *__restrict__ counters) {
    extern __shared__ float smem[];
    float *sq = smem;            // hd
    float *sc = smem + hd;       // scores of this split
    __shared__ float red[4];
    __shared__ int s_last;
    const int n_ctx = d_pos ? (*d_pos + 1) : nctx_imm;
    const int h = blockIdx.x, split = blockIdx.y, kvh = h / (nh / nkv);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_pages = (n_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const int pps = (n_pages + ATTN_NSPLIT - 1) / ATTN_NSPLIT;
    const int p0 = split * pps, p1 = min(n_pages, p0 + pps);
    const int t0 = p0 * PAGE_TOKENS, t1 = min(n_ctx, p1 * PAGE_TOKENS);
    const int nt = max(0, t1 - t0);
    const float scale = 1.0f / sqrtf((float)hd);
    for (int j = tid; j < hd; j += 128) sq[j] = q[(int64_t)h * hd + j];
    __syncthreads();
    // ---- phase 1: scores
    const int key = tid >> 3, sub = tid & 7;
    const int nchunk = hd >> 3;   // 16-byte chunks per key row
    float lmax = -INFINITY;
    for (int p = p0; p < p1; ++p) {
        const int t = p * PAGE_TOKENS + key;
        const __half *kr = kc + (((int64_t)page_table[p] * nkv + kvh) * PAGE_TOKENS + key) * hd;
        float dot = 0.f;
        if (t < n_ctx) {
            for (int c = sub; c < nchunk; c += 8) {
                uint4 raw = __ldg((const uint4 *)(kr + c * 8));
                const __half2 *h2 = (const __half2 *)&raw;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    float2 f = __half22float2(h2[e]);
                    dot = fmaf(sq[c * 8 + 2 * e], f.x, dot);
                    dot = fmaf(sq[c * 8 + 2 * e + 1], f.y, dot);
                }
            }
        }
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        dot += __shfl_xor_sync(0xffffffffu, dot, 4);
        if (t < n_ctx) {
            float s = dot * scale;
            if (sub == 0) sc[t - t0] = s;
            lmax = fmaxf(lmax, s);
        }
    }
    lmax = warp_max(lmax);
    if (lane == 0) red[warp] = lmax;
    __syncthreads();
    const float m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    __syncthreads();
    // ---- softmax numerators
    float lsum = 0.f;
    for (int i = tid; i < nt; i += 128) {
        float pexp = expf(sc[i] - m);
        sc[i] = pexp;
        lsum += pexp;
    }
    lsum = warp_sum(lsum);
    if (lane == 0) red[warp] = lsum;
    __syncthreads();
    const float l = red[0] + red[1] + red[2] + red[3];
    // ---- phase 2: weighted values
    float o0 = 0.f, o1 = 0.f;
    const int d0 = tid, d1 = tid + 128;
    for (int p = p0; p < p1; ++p) {
        const __half *vr = vc + ((int64_t)page_table[p] * nkv + kvh) * PAGE_TOKENS * hd;
        const int tb = p * PAGE_TOKENS;
        const int lim = min(PAGE_TOKENS, n_ctx - tb);
        for (int k = 0; k < lim; ++k) {
            float pe = sc[tb + k - t0];
            if (d0 < hd) o0 = fmaf(pe, __half2float(vr[k * hd + d0]), o0);
            if (d1 < hd) o1 = fmaf(pe, __half2float(vr[k * hd + d1]), o1);
        }
    }
    float *pp = part + ((int64_t)h * ATTN_NSPLIT + split) * (hd + 2);
    if (tid == 0) { pp[0] = m; pp[1] = l; }
    if (d0 < hd) pp[2 + d0] = o0;
    if (d1 < hd) pp[2 + d1] = o1;
    // ---- merge by the last split of this head
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        int ticket = atomicAdd(&counters[h], 1);
        s_last = (ticket == ATTN_NSPLIT - 1);
        if (s_last) counters[h] = 0;   // leave the counter ready for the next launch
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const float *ph = part + (int64_t)h * ATTN_NSPLIT * (hd + 2);
    float M = -INFINITY;
#pragma unroll
    for (int s = 0; s < ATTN_NSPLIT; ++s) M = fmaxf(M, ph[s * (hd + 2)]);
    float Lt = 0.f, a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int s = 0; s < ATTN_NSPLIT; ++s) {
        const float *ps = ph + s * (hd + 2);
        float ls = ps[1];
        if (ls > 0.f) {
            float wgt = expf(ps[0] - M);
            Lt = fmaf(ls, wgt, Lt);
            if (d0 < hd) a0 = fmaf(ps[2 + d0], wgt, a0);
            if (d1 < hd) a1 = fmaf(ps[2 + d1], wgt, a1);
        }
    }
    if (d0 < hd) out[(int64_t)h * hd + d0] = a0 / Lt;
    if (d1 < hd) out[(int64_t)h * hd + d1] = a1 / Lt;
}
void launch_attn_decode(const float *d_q, int nh, int nkv, int hd, const int *d_pos, int nctx_imm, const __half *d_k, const __half *d_v,
                        const int32_t *d_page_table, float *d_out, void *d_scratch, int max_ctx, cudaStream_t st) {
    float *part = (float *)d_scratch;
    int *counters = (int *)((char *)d_scratch + (uint64_t)nh * ATTN_NSPLIT * (hd + 2) * 4);
    int max_pages = (max_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    int pps = (max_pages + ATTN_NSPLIT - 1) / ATTN_NSPLIT;
    size_t smem = (size_t)(hd + pps * PAGE_TOKENS) * sizeof(float);
    k_attn_decode<<<dim3(nh, ATTN_NSPLIT), 128, smem, st>>>(d_q, nh, nkv, hd, d_pos, nctx_imm, d_k, d_v, d_page_table, d_out, part, counters);
}

// =============================================================================== greedy sampler
// Stage 1: ARGMAX_BLOCKS blocks scan strided slices (warp-shuffle arg-reduce, lowest index wins ties);
// the last block to finish (ticket counter) reduces the per-block partials.
uint64_t argmax_scratch_bytes() { return ARGMAX_BLOCKS * 8 + 4; }

__global__ void __launch_bounds__(256) k_argmax(const float *__restrict__ x, int n, int *__restrict__ out, float *__restrict__ pv,
                                                int *__restrict__ pi, int *__restrict__ counter) {
    __shared__ float sv[8];
    __shared__ int si[8];
    __shared__ int s_last;
    float v = -INFINITY;
    int idx = 0x7fffffff;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) arg_better(v, idx, x[i], i);
    warp_argmax(v, idx);
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = v; si[threadIdx.x >> 5] = idx; }
    __syncthreads();
    if (threadIdx.x < 32) {
        v = threadIdx.x < 8 ? sv[threadIdx.x] : -INFINITY;
        idx = threadIdx.x < 8 ? si[threadIdx.x] : 0x7fffffff;
        warp_argmax(v, idx);
        if (threadIdx.x == 0) {
            pv[blockIdx.x] = v;
            pi[blockIdx.x] = idx;
            __threadfence();
            int ticket = atomicAdd(counter, 1);
            s_last = (ticket == (int)gridDim.x - 1);
            if (s_last) *counter = 0;
        }
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x < 32) {
        v = -INFINITY;
        idx = 0x7fffffff;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += 32) arg_better(v, idx, pv[b], pi[b]);
        warp_argmax(v, idx);
        if (threadIdx.x == 0) *out = idx;
    }
}
__global__ void k_argmax_partials(const float *__restrict__ pv, const int *__restrict__ pi, int n, int *__restrict__ out) {
    float v = -INFINITY;
    int idx = 0x7fffffff;
    for (int b = threadIdx.x; b < n; b += 32) arg_better(v, idx, pv[b], pi[b]);
    warp_argmax(v, idx);
    if (threadIdx.x == 0) *out = idx;
}
void launch_argmax_partials(const float *pmax, const int *pidx, int n, int32_t *d_out, cudaStream_t st) {
    k_argmax_partials<<<1, 32, 0, st>>>(pmax, pidx, n, d_out);
}
void launch_argmax(const float *d_logits, int64_t n, int32_t *d_out, void *d_scratch, cudaStream_t st) {
    float *pv = (float *)d_scratch;
    int *pi = (int *)((char *)d_scratch + ARGMAX_BLOCKS * 4);
    int *counter = (int *)((char *)d_scratch + ARGMAX_BLOCKS * 8);
    k_argmax<<<ARGMAX_BLOCKS, 256, 0, st>>>(d_logits, (int)n, d_out, pv, pi, counter);
}
