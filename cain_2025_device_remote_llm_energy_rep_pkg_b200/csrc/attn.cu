// attn.cu — fused decode attention: RoPE + KV append + paged attention + split merge + activation grid
// (stages a3.4 and a3.5 of Ollama's /api/generate decode loop).
//
// grid (nh, ATTN_NSPLIT) launched as clusters (1, ATTN_NSPLIT, 1): the ATTN_NSPLIT CTAs of a cluster split the
// context of one query head by pages.  Every CTA rotates the query itself; the CTA whose range holds the
// newest position rotates/rounds the new key and value from the QKV output (and, for the first head of a
// KV group, appends them to the paged f16 cache).  Partials (max, sum, weighted values) stay in shared
// memory; after a cluster barrier the rank-0 CTA merges them through distributed shared memory and writes
// the head's output on the activation grid (the input format of the O-projection GEMV).
// Numerics follow oracle/model.py: fp32 scores/softmax over f16-rounded K and V (the newest token included).
#include "kernels.cuh"
#include "layout.cuh"

__device__ __forceinline__ void pdl_wait_a() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_a() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ float ld_dsmem_f32(const float *local_ptr, uint32_t rank) {
    uint32_t la = (uint32_t)__cvta_generic_to_shared(local_ptr), ra;
    float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(rank));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
    return v;
}

struct AttnArgs {
    const float *qkv;      // [nh*hd | nkv*hd | nkv*hd] fp32, un-rotated
    int nh, nkv, hd;
    const int *d_pos;
    int pos_imm;
    const float *cs, *sn;  // [max_ctx][hd/2]
    __half *kc, *vc;       // paged cache of this layer
    const int32_t *page_table;
    float *out;            // optional fp32 [nh*hd]
    uint8_t *act_out;      // optional A3 image of the output
    int pps_max;           // pages per split the shared-memory score buffer is sized for
};

__device__ __forceinline__ void l2_prefetch_a(const void *p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// NSPLIT = ATTN_NSPLIT: one cluster per head (long contexts).  NSPLIT = 1: one plain CTA per head (contexts up to
// ATTN_SHORT_CTX), no cluster barrier, no remote reads.
template <int NSPLIT>
__global__ void __launch_bounds__(128) k_attn_fused(const AttnArgs a) {
    extern __shared__ __align__(16) float sm[];
    const int hd = a.hd, half = hd >> 1;
    float *sq = sm;                                   // hd
    float *part = sq + hd;                            // hd + 2 : m, l, o[]   (read remotely by rank 0)
    float *og = part + hd + 2 + 2;                    // 1024 : per token-group partial outputs (16-byte aligned: hd+4 is a multiple of 4)
    float *sc = og + 1024;                            // pps_max * 16 scores
    __half *knew = (__half *)(sc + a.pps_max * PAGE_TOKENS);   // hd
    __half *vnew = knew + hd;                         // hd
    int *spage = (int *)(vnew + hd);                  // pps_max physical page ids of this CTA's range
    __shared__ float red[4];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int h = blockIdx.x, split = blockIdx.y, kvh = h / (a.nh / a.nkv);
    if (tid == 0) pdl_launch_a();
    // The position was written at the end of the previous token, the cache rows of older tokens long before: neither depends
    // on the kernel in front of this one, so the pages of this CTA's range are pulled into L2 while that kernel still runs.
    const int pos = a.d_pos ? __ldcg(a.d_pos) : a.pos_imm;
    const int n_ctx = pos + 1;
    const int n_pages = (n_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const int pps = (n_pages + NSPLIT - 1) / NSPLIT;
    const int p0 = split * pps, p1 = min(n_pages, p0 + pps);
    for (int p = p0 + tid; p < p1; p += 128) {
        const int pg = a.page_table[p];
        spage[p - p0] = pg;
        const size_t off = (((size_t)pg * a.nkv + kvh) * PAGE_TOKENS) * hd;
        l2_prefetch_a(a.kc + off, (uint32_t)(PAGE_TOKENS * hd * 2));
        l2_prefetch_a(a.vc + off, (uint32_t)(PAGE_TOKENS * hd * 2));
    }
    pdl_wait_a();
    const int t0 = p0 * PAGE_TOKENS, t1 = min(n_ctx, p1 * PAGE_TOKENS);
    const int nt = max(0, t1 - t0);
    const bool has_new = (pos >= t0 && pos < t1);
    const float scale = 1.0f / sqrtf((float)hd);
    // ---- RoPE on the query (every CTA), on the new key (the CTA that owns position pos)
    const float *cs = a.cs + (int64_t)pos * half, *sn = a.sn + (int64_t)pos * half;
    for (int j = tid; j < half; j += 128) {
        const float c = cs[j], s = sn[j];
        const float qa = __ldcg(a.qkv + (int64_t)h * hd + j), qb = __ldcg(a.qkv + (int64_t)h * hd + j + half);   // L2: written by the previous kernel
        sq[j] = __fsub_rn(__fmul_rn(qa, c), __fmul_rn(qb, s));
        sq[j + half] = __fadd_rn(__fmul_rn(qb, c), __fmul_rn(qa, s));
        if (has_new) {
            const float *kr = a.qkv + (int64_t)(a.nh + kvh) * hd, *vr = a.qkv + (int64_t)(a.nh + a.nkv + kvh) * hd;
            const float ka = __ldcg(kr + j), kb = __ldcg(kr + j + half);
            knew[j] = __float2half_rn(__fsub_rn(__fmul_rn(ka, c), __fmul_rn(kb, s)));
            knew[j + half] = __float2half_rn(__fadd_rn(__fmul_rn(kb, c), __fmul_rn(ka, s)));
            vnew[j] = __float2half_rn(__ldcg(vr + j));
            vnew[j + half] = __float2half_rn(__ldcg(vr + j + half));
        }
    }
    __syncthreads();
    if (has_new && (h % (a.nh / a.nkv)) == 0) {   // one writer per KV head appends to the cache
        const int page = spage[pos / PAGE_TOKENS - p0], slot = pos % PAGE_TOKENS;
        __half *kd = a.kc + (((int64_t)page * a.nkv + kvh) * PAGE_TOKENS + slot) * hd;
        __half *vd = a.vc + (((int64_t)page * a.nkv + kvh) * PAGE_TOKENS + slot) * hd;
        for (int j = tid; j < hd; j += 128) { kd[j] = knew[j]; vd[j] = vnew[j]; }
    }
    // ---- phase 1: scores.  8 lanes per key, 16 keys = one page per step; four pages are in flight per thread (all loads of
    //      the four pages are issued before the first dot product) so a split of a long context costs a few L2 round trips.
    const int key = tid >> 3, sub = tid & 7, nchunk = hd >> 3;   // 16-byte chunks per key row: 12 / 16 / 32
    float lmax = -INFINITY;
    for (int pb = p0; pb < p1; pb += 4) {
        uint4 kv[4][4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = pb + u, tk = p * PAGE_TOKENS + key;
            if (p < p1 && tk < n_ctx) {
                const __half *kr = (tk == pos) ? knew : a.kc + (((int64_t)spage[p - p0] * a.nkv + kvh) * PAGE_TOKENS + key) * hd;
#pragma unroll
                for (int c4 = 0; c4 < 4; ++c4) {
                    const int c = sub + 8 * c4;
                    if (c < nchunk) kv[u][c4] = *(const uint4 *)(kr + c * 8);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = pb + u, tk = p * PAGE_TOKENS + key;
            float dot = 0.f;
            if (p < p1 && tk < n_ctx) {
#pragma unroll
                for (int c4 = 0; c4 < 4; ++c4) {
                    const int c = sub + 8 * c4;
                    if (c < nchunk) {
                        const __half2 *h2 = (const __half2 *)&kv[u][c4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 f = __half22float2(h2[e]);
                            dot = fmaf(sq[c * 8 + 2 * e], f.x, dot);
                            dot = fmaf(sq[c * 8 + 2 * e + 1], f.y, dot);
                        }
                    }
                }
            }
            dot += __shfl_xor_sync(0xffffffffu, dot, 1);
            dot += __shfl_xor_sync(0xffffffffu, dot, 2);
            dot += __shfl_xor_sync(0xffffffffu, dot, 4);
            if (p < p1 && tk < n_ctx) {
                const float sv = dot * scale;
                if (sub == 0) sc[tk - t0] = sv;
                lmax = fmaxf(lmax, sv);
            }
        }
    }
    lmax = warp_max(lmax);
    if (lane == 0) red[warp] = lmax;
    __syncthreads();
    const float m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    __syncthreads();
    float lsum = 0.f;
    for (int i = tid; i < nt; i += 128) {
        const float pe = expf(sc[i] - m);
        sc[i] = pe;
        lsum += pe;
    }
    lsum = warp_sum(lsum);
    if (lane == 0) red[warp] = lsum;
    __syncthreads();
    const float l = red[0] + red[1] + red[2] + red[3];
    // ---- phase 2: weighted values.  A thread owns 8 dims (one 16-byte load per token); hd/8 threads cover a token and the
    //      128 threads work on 128/(hd/8) tokens at once; eight tokens per thread are in flight.
    const int tpt = hd >> 3;                      // threads per token: 12 / 16 / 32
    const int ngrp = 128 / tpt;                   // token groups: 10 / 8 / 4
    const int grp = tid / tpt, dq = tid - grp * tpt;
    float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (grp < ngrp) {
        for (int ib = grp; ib < nt; ib += 8 * ngrp) {
            uint4 vv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = ib + u * ngrp;
                if (i < nt) {
                    const int tk = t0 + i;
                    const __half *vr = (tk == pos) ? vnew
                                                   : a.vc + (((int64_t)spage[tk / PAGE_TOKENS - p0] * a.nkv + kvh) * PAGE_TOKENS + (tk % PAGE_TOKENS)) * hd;
                    vv[u] = *(const uint4 *)(vr + dq * 8);
                }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = ib + u * ngrp;
                if (i < nt) {
                    const float pe = sc[i];
                    const __half2 *h2 = (const __half2 *)&vv[u];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 f = __half22float2(h2[e]);
                        o[2 * e] = fmaf(pe, f.x, o[2 * e]);
                        o[2 * e + 1] = fmaf(pe, f.y, o[2 * e + 1]);
                    }
                }
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) og[grp * hd + dq * 8 + e] = o[e];
    }
    __syncthreads();
    for (int d = tid; d < hd; d += 128) {
        float acc = 0.f;
        for (int gq = 0; gq < ngrp; ++gq) acc += og[gq * hd + d];   // fixed order
        part[2 + d] = acc;
    }
    if (tid == 0) { part[0] = m; part[1] = l; }
    // ---- merge across the cluster through distributed shared memory (NSPLIT > 1), or finish locally
    if (NSPLIT > 1) cluster_sync_all();
    else __syncthreads();
    if (NSPLIT == 1 || cluster_rank() == 0) {
        if (NSPLIT > 1) {
            float M = -INFINITY;
#pragma unroll
            for (int r = 0; r < NSPLIT; ++r) M = fmaxf(M, ld_dsmem_f32(part, r));
            float wgt[NSPLIT], L = 0.f;
#pragma unroll
            for (int r = 0; r < NSPLIT; ++r) {
                const float lr = ld_dsmem_f32(part + 1, r);
                wgt[r] = lr > 0.f ? expf(ld_dsmem_f32(part, r) - M) : 0.f;
                L = fmaf(lr, wgt[r], L);
            }
            for (int d = tid; d < hd; d += 128) {
                float acc = 0.f;
#pragma unroll
                for (int r = 0; r < NSPLIT; ++r)
                    if (wgt[r] > 0.f) acc = fmaf(ld_dsmem_f32(part + 2 + d, r), wgt[r], acc);
                const float ov = acc / L;
                sq[d] = ov;                                   // the query is no longer needed
                if (a.out) a.out[(int64_t)h * hd + d] = ov;
            }
        } else {
            for (int d = tid; d < hd; d += 128) {
                const float ov = part[2 + d] / l;             // same rounding as the merge of a single split (weight exp(0) = 1)
                sq[d] = ov;
                if (a.out) a.out[(int64_t)h * hd + d] = ov;
            }
        }
        __syncthreads();
        if (a.act_out && warp < hd / QK) {
            float s;
            const int X = act_grid_lane(sq[warp * QK + lane], s);
            a3_store_block(a.act_out, (int64_t)a.nh * hd, (int64_t)h * (hd / QK) + warp, lane, X, s);
        }
        if (a.act_out && hd / QK > 4)                     // hd 256: 8 blocks, 4 warps
            for (int b = warp + 4; b < hd / QK; b += 4) {
                float s;
                const int X = act_grid_lane(sq[b * QK + lane], s);
                a3_store_block(a.act_out, (int64_t)a.nh * hd, (int64_t)h * (hd / QK) + b, lane, X, s);
            }
    }
    if (NSPLIT > 1) cluster_sync_all();   // the other CTAs' shared memory must outlive rank 0's remote reads
}

static size_t attn_fused_smem(int hd, int pps) {
    return (size_t)(hd + (hd + 4) + 1024 + pps * PAGE_TOKENS) * 4 + (size_t)2 * hd * 2 + (size_t)pps * 4 + 16;
}

// max_ctx sizes the score buffer.  Contexts up to ATTN_SHORT_CTX run one CTA per head; longer ones one cluster per head.
cudaError_t launch_attn_fused(const float *d_qkv, int nh, int nkv, int hd, const int *d_pos, int pos_imm, const float *d_cos,
                              const float *d_sin, __half *d_k, __half *d_v, const int32_t *d_page_table, float *d_out, void *d_act_out,
                              int max_ctx, cudaStream_t st, bool pdl) {
    AttnArgs a;
    a.qkv = d_qkv; a.nh = nh; a.nkv = nkv; a.hd = hd; a.d_pos = d_pos; a.pos_imm = pos_imm; a.cs = d_cos; a.sn = d_sin;
    a.kc = d_k; a.vc = d_v; a.page_table = d_page_table; a.out = d_out; a.act_out = (uint8_t *)d_act_out;
    const bool one = max_ctx <= ATTN_SHORT_CTX;
    const int nsplit = one ? 1 : ATTN_NSPLIT;
    int max_pages = (max_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    a.pps_max = (max_pages + nsplit - 1) / nsplit;
    size_t smem = attn_fused_smem(hd, a.pps_max);
    if (smem > 48 * 1024) {
        cudaError_t e = one ? cudaFuncSetAttribute(k_attn_fused<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                            : cudaFuncSetAttribute(k_attn_fused<ATTN_NSPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nh, nsplit);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (!one) {
        at[na].id = cudaLaunchAttributeClusterDimension;
        at[na].val.clusterDim.x = 1;
        at[na].val.clusterDim.y = ATTN_NSPLIT;
        at[na].val.clusterDim.z = 1;
        ++na;
    }
    if (pdl) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = at;
    cfg.numAttrs = na;
    return one ? cudaLaunchKernelEx(&cfg, k_attn_fused<1>, a) : cudaLaunchKernelEx(&cfg, k_attn_fused<ATTN_NSPLIT>, a);
}
