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
#include "attn_dev.cuh"
#include <stdlib.h>
#define ATTN_PRE_DEFAULT 0

#ifdef GEMV_EXP_NOWAIT
__device__ __forceinline__ void pdl_wait_a() {}
#else
__device__ __forceinline__ void pdl_wait_a() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif
__device__ __forceinline__ void pdl_launch_a() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// NSPLIT = ATTN_NSPLIT: one cluster per head (long contexts).  NSPLIT = 1: one plain CTA per head (contexts up to
// ATTN_SHORT_CTX), no cluster barrier, no remote reads.  The work itself is attn_dev.cuh.
template <int NSPLIT, int PRE>
__global__ void __launch_bounds__(128) k_attn_fused(const AttnArgs a) {
    extern __shared__ __align__(16) float sm[];
    const AttnSmem s = attn_carve(sm, a.hd, a.pps_max);
    const int tid = threadIdx.x, h = blockIdx.x, split = blockIdx.y;
    if (tid == 0) pdl_launch_a();
    const int pos = a.d_pos ? __ldcg(a.d_pos) : a.pos_imm;
    attn_pages(a, s, tid, h, split, NSPLIT, pos);
    AttnRegs R;
    if (PRE == 1) attn_preload<false>(a, s, tid, h, split, NSPLIT, pos, 0, R);
    pdl_wait_a();
    if (PRE == 2) attn_preload<true>(a, s, tid, h, split, NSPLIT, pos, 0, R);
    attn_main<NSPLIT == 1 ? MERGE_NONE : MERGE_CLUSTER, PRE>(a, s, tid, h, split, NSPLIT, pos, 0, R);
}

// One CTA of NG groups of 128 threads per head: the groups split the context as the CTAs of a cluster do, but merge through the
// CTA's shared memory (one __syncthreads instead of two cluster barriers and remote reads), and a layer's attention occupies nh SMs
// instead of all of them, so that the O projection's CTAs become resident (and request their weights) while attention runs.
template <int NG>
__global__ void __launch_bounds__(128 * NG, 1) k_attn_fused_cta(const AttnArgs a) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x & 127, grp = threadIdx.x >> 7, h = blockIdx.x;
    const AttnSmem s = attn_carve(sm + (size_t)grp * a.group_stride, a.hd, a.pps_max);
    if (threadIdx.x == 0) pdl_launch_a();
    const int pos = a.d_pos ? __ldcg(a.d_pos) : a.pos_imm;
    attn_pages(a, s, tid, h, grp, NG, pos);
    AttnRegs R;
    pdl_wait_a();
    attn_main<MERGE_CTA, 0, 2, 4>(a, s, tid, h, grp, NG, pos, grp + 1, R);   // named barrier grp + 1 (0 is __syncthreads)
}

static size_t attn_fused_smem(int hd, int pps) {
    return (size_t)attn_group_floats(hd, pps) * 4;
}

// max_ctx sizes the score buffer and picks the kernel: contexts up to ATTN_SHORT_CTX run one 128-thread CTA per head, up to ATTN_CTA_CTX one
// 1024-thread CTA per head, longer ones one cluster per head.
cudaError_t launch_attn_fused(const float *d_qkv, int nh, int nkv, int hd, const int *d_pos, int pos_imm, const float *d_cos,
                              const float *d_sin, __half *d_k, __half *d_v, const int32_t *d_page_table, float *d_out, void *d_act_out,
                              int max_ctx, cudaStream_t st, bool pdl) {
    AttnArgs a = {};
    a.qkv = d_qkv; a.nh = nh; a.nkv = nkv; a.hd = hd; a.d_pos = d_pos; a.pos_imm = pos_imm; a.cs = d_cos; a.sn = d_sin;
    a.kc = d_k; a.vc = d_v; a.page_table = d_page_table; a.out = d_out; a.act_out = (uint8_t *)d_act_out;
    const bool one = max_ctx <= ATTN_SHORT_CTX;
    const int nsplit = one ? 1 : ATTN_NSPLIT;
    int max_pages = (max_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    a.pps_max = (max_pages + nsplit - 1) / nsplit;
    size_t smem = attn_fused_smem(hd, a.pps_max);
    // B2_ATTN_PRE: 0 = loads in dependency order (q, then K, then V); 1 = K/V of older tokens fetched into registers BEFORE the
    // dependency wait (measured slower: the QKV GEMV owns every SM until it exits, there is no window before the wait);
    // 2 = after the wait, q first and then all of K and V in one burst, so the three round trips overlap.
    static int pre = -1;
    if (pre < 0) { const char *e = getenv("B2_ATTN_PRE"); pre = e ? atoi(e) : ATTN_PRE_DEFAULT; if (pre < 0 || pre > 2) pre = 0; }
    static int cta_merge = -1;   // contexts in (ATTN_SHORT_CTX, ATTN_CTA_CTX] run one 1024-thread CTA per head (k_attn_fused_cta); B2_ATTN_CTA=0: the cluster kernel
    if (cta_merge < 0) { const char *e = getenv("B2_ATTN_CTA"); cta_merge = (e && e[0] == '0') ? 0 : 1; }
    if (!one && cta_merge && max_ctx <= ATTN_CTA_CTX) {
        a.group_stride = attn_group_floats(hd, a.pps_max);
        const size_t smem_c = (size_t)ATTN_NSPLIT * a.group_stride * 4;
        if (smem_c > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(k_attn_fused_cta<ATTN_NSPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_c);
            if (e != cudaSuccess) return e;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(nh);
        cfg.blockDim = dim3(128 * ATTN_NSPLIT);
        cfg.dynamicSmemBytes = smem_c;
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = pdl ? 1 : 0;
        return cudaLaunchKernelEx(&cfg, k_attn_fused_cta<ATTN_NSPLIT>, a);
    }
    const void *fn = nullptr;
    if (one) fn = pre == 2 ? (const void *)k_attn_fused<1, 2> : pre == 1 ? (const void *)k_attn_fused<1, 1> : (const void *)k_attn_fused<1, 0>;
    else fn = pre == 2 ? (const void *)k_attn_fused<ATTN_NSPLIT, 2> : pre == 1 ? (const void *)k_attn_fused<ATTN_NSPLIT, 1> : (const void *)k_attn_fused<ATTN_NSPLIT, 0>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
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
    void *args[1] = {(void *)&a};
    return cudaLaunchKernelExC(&cfg, fn, args);
}
