// attn_dev.cuh — device side of the fused decode attention (see attn.cu for the design): one (head, context split) work item
// executed by a group of 128 threads (attn.cu: the group is the CTA, splits merge through distributed shared memory of a cluster).
// Numerics follow oracle/model.py: fp32 scores/softmax over f16-rounded K and V (the newest token included).
#pragma once
#include "kernels.cuh"
#include "layout.cuh"

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
    int group_stride;      // MERGE_CTA: floats of shared memory between the areas of two groups (splits) of one CTA
};
enum { MERGE_NONE = 0, MERGE_CLUSTER = 1, MERGE_CTA = 2 };   // splits of a head: none / CTAs of a cluster (DSMEM) / 128-thread groups of one CTA

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
__device__ __forceinline__ void l2_prefetch_a(const void *p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
// barrier of one 128-thread group (id 0 in a 128-thread CTA is __syncthreads)
__device__ __forceinline__ void group_sync(int bar_id) { asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory"); }

// floats of shared memory one group needs (16-byte multiple)
__host__ __device__ inline int attn_group_floats(int hd, int pps) {
    // sq hd | part hd+4 | og 1024 | sc pps*16 | knew,vnew hd halfs each = hd floats | spage pps | red 8
    return ((hd + (hd + 4) + 1024 + pps * PAGE_TOKENS + hd + pps + 8) + 3) & ~3;
}

struct AttnSmem {
    float *sq, *part, *og, *sc, *red;
    __half *knew, *vnew;
    int *spage;
};
__device__ __forceinline__ AttnSmem attn_carve(float *sm, int hd, int pps_max) {
    AttnSmem s;
    s.sq = sm;                                    // hd
    s.part = s.sq + hd;                           // hd + 2 (+2 pad): m, l, o[]   (read remotely by rank 0 under MERGE_CLUSTER)
    s.og = s.part + hd + 4;                       // 1024 : per token-group partial outputs (16-byte aligned: hd+4 is a multiple of 4)
    s.sc = s.og + 1024;                           // pps_max * 16 scores
    s.knew = (__half *)(s.sc + pps_max * PAGE_TOKENS);
    s.vnew = s.knew + hd;
    s.spage = (int *)(s.vnew + hd);               // pps_max physical page ids of this item's range
    s.red = (float *)(s.spage + pps_max);         // 8
    return s;
}

// Part 1 of an item, independent of the kernel/phase in front of it: the position was written at the end of the previous
// token and the cache rows of older tokens long before, so the page ids go to shared memory and the pages of this item's range
// are pulled into L2 while the producer of qkv still runs.
__device__ __forceinline__ void attn_pages(const AttnArgs &a, const AttnSmem &s, int tid, int h, int split, int nsplit, int pos) {
    const int kvh = h / (a.nh / a.nkv);
    const int n_pages = (pos + 1 + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const int pps = (n_pages + nsplit - 1) / nsplit;
    const int p0 = split * pps, p1 = min(n_pages, p0 + pps);
    for (int p = p0 + tid; p < p1; p += 128) {
        const int pg = a.page_table[p];
        s.spage[p - p0] = pg;
        const size_t off = (((size_t)pg * a.nkv + kvh) * PAGE_TOKENS) * a.hd;
        l2_prefetch_a(a.kc + off, (uint32_t)(PAGE_TOKENS * a.hd * 2));
        l2_prefetch_a(a.vc + off, (uint32_t)(PAGE_TOKENS * a.hd * 2));
    }
}

// What an item can fetch before the producer of qkv has finished (stand-alone kernel: before griddepcontrol.wait): the RoPE
// table row of this position, and the cached K rows (first four pages of the item's range) and V rows (first eight tokens per
// thread) of OLDER tokens.  With contexts up to 64 positions per split that is everything but q and the newest k/v, so the
// critical path after the wait is one L2 round trip for q plus arithmetic, not three dependent round trips.
struct AttnRegs {
    float c, s;        // cos/sin of this thread's rotation pair (threads < hd/2)
    float qa, qb, ka, kb, va, vb;   // q, and the newest k and v, of that pair (AFTER_WAIT only: they come from the QKV kernel)
    uint4 kv[4][4];    // phase 1, first batch: [page][chunk]
    uint4 vv[8];       // phase 2, first batch
};
template <bool AFTER_WAIT>
__device__ __forceinline__ void attn_preload(const AttnArgs &a, const AttnSmem &s, int tid, int h, int split, int nsplit, int pos, int bar_id, AttnRegs &R) {
    const int hd = a.hd, half = hd >> 1;
    const int kvh = h / (a.nh / a.nkv);
    const int n_ctx = pos + 1;
    const int n_pages = (n_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const int pps = (n_pages + nsplit - 1) / nsplit;
    const int p0 = split * pps, p1 = min(n_pages, p0 + pps);
    const int t0 = p0 * PAGE_TOKENS, t1 = min(n_ctx, p1 * PAGE_TOKENS);
    const int nt = max(0, t1 - t0);
    R.c = 1.f; R.s = 0.f;
    R.qa = R.qb = R.ka = R.kb = R.va = R.vb = 0.f;
    if (tid < half) {   // the loops below give pair j to thread j % 128: hd <= 256
        if (AFTER_WAIT) {   // q first: everything after it in this SM's load queue returns after it
            const bool has_new = pos >= t0 && pos < t1;
            R.qa = __ldcg(a.qkv + (int64_t)h * hd + tid);
            R.qb = __ldcg(a.qkv + (int64_t)h * hd + tid + half);
            if (has_new) {
                const float *kr = a.qkv + (int64_t)(a.nh + kvh) * hd, *vr = a.qkv + (int64_t)(a.nh + a.nkv + kvh) * hd;
                R.ka = __ldcg(kr + tid); R.kb = __ldcg(kr + tid + half);
                R.va = __ldcg(vr + tid); R.vb = __ldcg(vr + tid + half);
            }
        }
        R.c = __ldg(a.cs + (int64_t)pos * half + tid);
        R.s = __ldg(a.sn + (int64_t)pos * half + tid);
    }
    group_sync(bar_id);   // page ids (attn_pages) are in shared memory
    const int key = tid >> 3, sub = tid & 7, nchunk = hd >> 3;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int p = p0 + u, tk = p * PAGE_TOKENS + key;
        const bool ld = p < p1 && tk < n_ctx && tk != pos;
        const __half *kr = a.kc + (((int64_t)(ld ? s.spage[p - p0] : 0) * a.nkv + kvh) * PAGE_TOKENS + key) * hd;
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4) {
            const int c = sub + 8 * c4;
            R.kv[u][c4] = make_uint4(0u, 0u, 0u, 0u);
            if (ld && c < nchunk) R.kv[u][c4] = __ldcg((const uint4 *)(kr + c * 8));
        }
    }
    const int tpt = hd >> 3, ngrp = 128 / tpt;
    const int grp = tid / tpt, dq = tid - grp * tpt;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const int i = grp + u * ngrp, tk = t0 + i;
        R.vv[u] = make_uint4(0u, 0u, 0u, 0u);
        if (grp < ngrp && i < nt && tk != pos)
            R.vv[u] = __ldcg((const uint4 *)(a.vc + (((int64_t)s.spage[tk / PAGE_TOKENS - p0] * a.nkv + kvh) * PAGE_TOKENS + (tk % PAGE_TOKENS)) * hd + dq * 8));
    }
}

// Part 2: RoPE on q (and on the new k), KV append, scores, softmax, weighted values, merge of the splits, activation grid.
// All 128 threads of the group call it with the same (h, split, nsplit, pos).
// PRE: the first batches of K and V and the RoPE row were fetched by attn_preload into R.
// KD / VD: pages of keys / tokens of values (per thread) requested before the first of them is used
template <int MERGE, int PRE, int KD = 4, int VD = 8>   // PRE: 0 no preload, 1 attn_preload<false> before the wait, 2 attn_preload<true> after it (q, k, v in R too)
__device__ __forceinline__ void attn_main(const AttnArgs &a, const AttnSmem &s, int tid, int h, int split, int nsplit, int pos, int bar_id, AttnRegs &R) {
    const int hd = a.hd, half = hd >> 1;
    float *sq = s.sq, *part = s.part, *og = s.og, *sc = s.sc, *red = s.red;
    __half *knew = s.knew, *vnew = s.vnew;
    const int *spage = s.spage;
    const int lane = tid & 31, warp = tid >> 5;
    const int kvh = h / (a.nh / a.nkv);
    const int n_ctx = pos + 1;
    const int n_pages = (n_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const int pps = (n_pages + nsplit - 1) / nsplit;
    const int p0 = split * pps, p1 = min(n_pages, p0 + pps);
    const int t0 = p0 * PAGE_TOKENS, t1 = min(n_ctx, p1 * PAGE_TOKENS);
    const int nt = max(0, t1 - t0);
    const bool has_new = (pos >= t0 && pos < t1);
    const float scale = 1.0f / sqrtf((float)hd);
    // ---- RoPE on the query (every split), on the new key (the split that owns position pos)
    const float *cs = a.cs + (int64_t)pos * half, *sn = a.sn + (int64_t)pos * half;
    for (int j = tid; j < half; j += 128) {
        const float c = PRE != 0 ? R.c : cs[j], sv = PRE != 0 ? R.s : sn[j];
        const float qa = PRE == 2 ? R.qa : __ldcg(a.qkv + (int64_t)h * hd + j), qb = PRE == 2 ? R.qb : __ldcg(a.qkv + (int64_t)h * hd + j + half);   // L2: written by the previous kernel/phase
        sq[j] = __fsub_rn(__fmul_rn(qa, c), __fmul_rn(qb, sv));
        sq[j + half] = __fadd_rn(__fmul_rn(qb, c), __fmul_rn(qa, sv));
        if (has_new) {
            const float *kr = a.qkv + (int64_t)(a.nh + kvh) * hd, *vr = a.qkv + (int64_t)(a.nh + a.nkv + kvh) * hd;
            const float ka = PRE == 2 ? R.ka : __ldcg(kr + j), kb = PRE == 2 ? R.kb : __ldcg(kr + j + half);
            knew[j] = __float2half_rn(__fsub_rn(__fmul_rn(ka, c), __fmul_rn(kb, sv)));
            knew[j + half] = __float2half_rn(__fadd_rn(__fmul_rn(kb, c), __fmul_rn(ka, sv)));
            vnew[j] = __float2half_rn(PRE == 2 ? R.va : __ldcg(vr + j));
            vnew[j + half] = __float2half_rn(PRE == 2 ? R.vb : __ldcg(vr + j + half));
        }
    }
    group_sync(bar_id);
    if (has_new && (h % (a.nh / a.nkv)) == 0) {   // one writer per KV head appends to the cache
        const int page = spage[pos / PAGE_TOKENS - p0], slot = pos % PAGE_TOKENS;
        __half *kd = a.kc + (((int64_t)page * a.nkv + kvh) * PAGE_TOKENS + slot) * hd;
        __half *vd = a.vc + (((int64_t)page * a.nkv + kvh) * PAGE_TOKENS + slot) * hd;
        for (int j = tid; j < hd; j += 128) { kd[j] = knew[j]; vd[j] = vnew[j]; }
    }
    // ---- phase 1: scores.  8 lanes per key, 16 keys = one page per step; four pages are in flight per thread (all loads of
    //      the four pages are issued before the first dot product) so a split of a long context costs a few L2 round trips.
    //      Cache rows are read with ld.global.cg: in the persistent kernel this SM's L1 may hold a line of an earlier token.
    const int key = tid >> 3, sub = tid & 7, nchunk = hd >> 3;   // 16-byte chunks per key row: 12 / 16 / 32
    float lmax = -INFINITY;
    for (int pb = p0; pb < p1; pb += KD) {
        uint4 kv[KD][4];
#pragma unroll
        for (int u = 0; u < KD; ++u) {
            const int p = pb + u, tk = p * PAGE_TOKENS + key;
            const bool valid = p < p1 && tk < n_ctx, fresh = tk == pos;
            const bool pre = PRE != 0 && pb == p0;   // older tokens of the first batch are already in registers
            const __half *kr = fresh ? knew : a.kc + (((int64_t)((valid && !pre) ? spage[p - p0] : 0) * a.nkv + kvh) * PAGE_TOKENS + key) * hd;
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) {
                const int c = sub + 8 * c4;
                kv[u][c4] = (pre && u < 4) ? R.kv[u < 4 ? u : 0][c4] : make_uint4(0u, 0u, 0u, 0u);   // defined on every path (see the note on buf0..3 in gemv_dev.cuh)
                if (valid && c < nchunk) {
                    if (fresh) kv[u][c4] = *(const uint4 *)(kr + c * 8);
                    else if (!pre) kv[u][c4] = __ldcg((const uint4 *)(kr + c * 8));
                }
            }
        }
#pragma unroll
        for (int u = 0; u < KD; ++u) {
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
    group_sync(bar_id);
    const float m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    group_sync(bar_id);
    float lsum = 0.f;
    for (int i = tid; i < nt; i += 128) {
        const float pe = expf(sc[i] - m);
        sc[i] = pe;
        lsum += pe;
    }
    lsum = warp_sum(lsum);
    if (lane == 0) red[warp] = lsum;
    group_sync(bar_id);
    const float l = red[0] + red[1] + red[2] + red[3];
    // ---- phase 2: weighted values.  A thread owns 8 dims (one 16-byte load per token); hd/8 threads cover a token and the
    //      128 threads work on 128/(hd/8) tokens at once; eight tokens per thread are in flight.
    const int tpt = hd >> 3;                      // threads per token: 12 / 16 / 32
    const int ngrp = 128 / tpt;                   // token groups: 10 / 8 / 4
    const int grp = tid / tpt, dq = tid - grp * tpt;
    float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (grp < ngrp) {
        for (int ib = grp; ib < nt; ib += VD * ngrp) {
            uint4 vv[VD];
#pragma unroll
            for (int u = 0; u < VD; ++u) {
                const int i = ib + u * ngrp;
                const bool pre = PRE != 0 && ib == grp;
                vv[u] = pre ? R.vv[u < 8 ? u : 0] : make_uint4(0u, 0u, 0u, 0u);
                if (i < nt) {
                    const int tk = t0 + i;
                    if (tk == pos) vv[u] = *(const uint4 *)(vnew + dq * 8);
                    else if (!pre)
                        vv[u] = __ldcg((const uint4 *)(a.vc + (((int64_t)spage[tk / PAGE_TOKENS - p0] * a.nkv + kvh) * PAGE_TOKENS + (tk % PAGE_TOKENS)) * hd + dq * 8));
                }
            }
#pragma unroll
            for (int u = 0; u < VD; ++u) {
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
    group_sync(bar_id);
    for (int d = tid; d < hd; d += 128) {
        float acc = 0.f;
        for (int gq = 0; gq < ngrp; ++gq) acc += og[gq * hd + d];   // fixed order
        part[2 + d] = acc;
    }
    if (tid == 0) { part[0] = m; part[1] = l; }
    // ---- merge of the splits (fixed order r = 0..nsplit-1 whichever way the partials travel), or finish locally
    bool finisher = true;
    if (MERGE == MERGE_CLUSTER) {
        cluster_sync_all();
        finisher = cluster_rank() == 0;
    } else if (MERGE == MERGE_CTA) {
        __syncthreads();       // every group of the CTA has left its partial (m, l, o[]) in its own shared-memory area
        finisher = split == 0;
    } else {
        group_sync(bar_id);
    }
    if (finisher) {
        if (MERGE != MERGE_NONE && nsplit > 1) {
            auto pm = [&](int r, int idx) -> float {
                return MERGE == MERGE_CLUSTER ? ld_dsmem_f32(part + idx, r) : (part + (r - split) * a.group_stride)[idx];
            };
            float M = -INFINITY;
#pragma unroll
            for (int r = 0; r < ATTN_NSPLIT; ++r)
                if (r < nsplit) M = fmaxf(M, pm(r, 0));
            float wgt[ATTN_NSPLIT], L = 0.f;
#pragma unroll
            for (int r = 0; r < ATTN_NSPLIT; ++r) {
                wgt[r] = 0.f;
                if (r < nsplit) {
                    const float lr = pm(r, 1);
                    wgt[r] = lr > 0.f ? expf(pm(r, 0) - M) : 0.f;
                    L = fmaf(lr, wgt[r], L);
                }
            }
            for (int d = tid; d < hd; d += 128) {
                float acc = 0.f;
#pragma unroll
                for (int r = 0; r < ATTN_NSPLIT; ++r)
                    if (wgt[r] > 0.f) acc = fmaf(pm(r, 2 + d), wgt[r], acc);
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
        group_sync(bar_id);
        if (a.act_out)
            for (int b = warp; b < hd / QK; b += 4) {         // hd 256: 8 blocks, 4 warps
                float sv;
                const int X = act_grid_lane(sq[b * QK + lane], sv);
                a3_store_block(a.act_out, (int64_t)a.nh * hd, (int64_t)h * (hd / QK) + b, lane, X, sv);
            }
    }
    if (MERGE == MERGE_CLUSTER) cluster_sync_all();   // the other CTAs' shared memory must outlive rank 0's remote reads
}
