// sampler.cu — top-k / temperature / top-p sampler (Ollama's default sampling family).
// Stage 1: SAMPLE_BLOCKS blocks each select the top-k of their slice in shared memory (k rounds of a
// warp-shuffle arg-reduce).  Stage 2: the last block to finish merges the SAMPLE_BLOCKS*k candidates,
// then one warp applies temperature softmax, the top-p cut and the inverse-CDF draw with uniform u.
// Contract (mirrored by oracle/sampler.py): candidates sorted by (logit desc, index asc);
// p = softmax(logit/T) over the k candidates; keep the shortest prefix with cumulative p >= top_p;
// pick the first candidate whose cumulative p >= u * (kept mass).
// Repetition penalty (Ollama's options.repeat_penalty / repeat_last_n, llama.cpp's rule): before the selection, the logit of every
// distinct token among the last `last_n` tokens of the context (prompt included) is divided by the penalty when positive and
// multiplied by it otherwise; penalty 1 (or 0) switches it off.
#include "kernels.cuh"
#include "sampler.cuh"

__device__ __forceinline__ void sel_better(float &v, int &i, float ov, int oi) {
    if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}
__device__ void block_argmax_256(float &v, int &i, float *sv, int *si) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, v, o);
        int oi = __shfl_xor_sync(0xffffffffu, i, o);
        sel_better(v, i, ov, oi);
    }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = v; si[threadIdx.x >> 5] = i; }
    __syncthreads();
    v = sv[0]; i = si[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) sel_better(v, i, sv[w], si[w]);
    __syncthreads();
}

__global__ void __launch_bounds__(256) k_sample_topk(const float *__restrict__ logits, int n, int k, float temperature, float top_p,
                                                     const float *__restrict__ d_u, float u_imm, int *__restrict__ out, float *cv, int *ci,
                                                     int *counter, const int *__restrict__ ctx, const int *__restrict__ d_ctx_len, int ctx_len_imm,
                                                     float penalty, int last_n) {
    extern __shared__ float sm[];
    __shared__ float sv[8];
    __shared__ int si[8];
    __shared__ int s_last;
    const int tid = threadIdx.x;
    const int per = (n + gridDim.x - 1) / gridDim.x;
    const int i0 = blockIdx.x * per, i1 = min(n, i0 + per);
    const int cnt = max(0, i1 - i0);
    float *vals = sm;   // slice copy, selected entries are overwritten with -inf
    for (int j = tid; j < cnt; j += 256) vals[j] = logits[i0 + j];
    __syncthreads();
    if (ctx && penalty > 0.f && penalty != 1.0f && last_n > 0) {
        const int len = d_ctx_len ? *d_ctx_len : ctx_len_imm, w0 = max(0, len - last_n);
        for (int j = w0 + tid; j < len; j += 256) {
            const int t = ctx[j];
            if (t >= i0 && t < i1) {
                bool first = true;   // a token that occurs several times in the window is penalised once
                for (int q = w0; q < j; ++q) first = first && ctx[q] != t;
                if (first) { const float v = vals[t - i0]; vals[t - i0] = v > 0.f ? __fdiv_rn(v, penalty) : __fmul_rn(v, penalty); }
            }
        }
        __syncthreads();
    }
    for (int r = 0; r < k; ++r) {
        float v = -INFINITY;
        int idx = 0x7fffffff;
        for (int j = tid; j < cnt; j += 256) sel_better(v, idx, vals[j], i0 + j);
        block_argmax_256(v, idx, sv, si);
        if (tid == 0) {
            cv[blockIdx.x * k + r] = v;
            ci[blockIdx.x * k + r] = idx;
            if (idx != 0x7fffffff) vals[idx - i0] = -INFINITY;
        }
        __syncthreads();
    }
    if (tid == 0) {
        __threadfence();
        int ticket = atomicAdd(counter, 1);
        s_last = (ticket == (int)gridDim.x - 1);
        if (s_last) *counter = 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // ---- merge: candidates into shared memory (values then indices)
    const int nc = gridDim.x * k;
    float *mv = sm;
    int *mi = (int *)(sm + nc);
    float *tv = (float *)(mi + nc);   // k selected values
    int *ti = (int *)(tv + k);
    __syncthreads();
    for (int j = tid; j < nc; j += 256) { mv[j] = cv[j]; mi[j] = ci[j]; }
    __syncthreads();
    for (int r = 0; r < k; ++r) {
        float v = -INFINITY;
        int idx = 0x7fffffff, slot = -1;
        for (int j = tid; j < nc; j += 256) {
            float ov = mv[j];
            int oi = mi[j];
            if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; slot = j; }
        }
        // arg-reduce carrying the slot: pack (idx) comparisons, then the winner clears its slot
        float bv = v;
        int bi = idx;
        block_argmax_256(bv, bi, sv, si);
        if (slot >= 0 && idx == bi && v == bv) mv[slot] = -INFINITY, mi[slot] = 0x7fffffff;
        if (tid == 0) { tv[r] = bv; ti[r] = bi; }
        __syncthreads();
    }
    if (tid == 0) {
        // temperature softmax over the k candidates (sorted descending), top-p cut, inverse-CDF draw
        int kk = 0;
        while (kk < k && ti[kk] != 0x7fffffff) ++kk;
        float mx = tv[0], sum = 0.f;
        for (int j = 0; j < kk; ++j) { tv[j] = expf((tv[j] - mx) / temperature); sum += tv[j]; }
        float cum = 0.f;
        int keep = kk;
        for (int j = 0; j < kk; ++j) {
            cum += tv[j] / sum;
            if (cum >= top_p) { keep = j + 1; break; }
        }
        float mass = 0.f;
        for (int j = 0; j < keep; ++j) mass += tv[j];
        float u = d_u ? *d_u : u_imm;
        float target = u * mass, run = 0.f;
        int pick = ti[keep - 1];
        for (int j = 0; j < keep; ++j) {
            run += tv[j];
            if (run >= target) { pick = ti[j]; break; }
        }
        *out = pick;
    }
}

uint64_t sample_scratch_bytes(int k) { return (uint64_t)SAMPLE_BLOCKS * k * 8 + 4; }

int launch_sample_topk(const float *d_logits, int64_t n, int k, float temperature, float top_p, const float *d_u, float u_imm,
                       int32_t *d_out, void *d_scratch, cudaStream_t st, const int *d_ctx, const int *d_ctx_len, int ctx_len_imm, float penalty,
                       int last_n) {
    if (k < 1 || k > SAMPLE_MAX_K) { b2_set_error("sampler: top_k %d out of range [1,%d]", k, SAMPLE_MAX_K); return B2_EINVAL; }
    if (!(temperature > 0.f)) { b2_set_error("sampler: temperature must be > 0"); return B2_EINVAL; }
    int per = (int)((n + SAMPLE_BLOCKS - 1) / SAMPLE_BLOCKS);
    size_t smem1 = (size_t)per * 4;
    size_t smem2 = (size_t)SAMPLE_BLOCKS * k * 8 + (size_t)k * 8;
    size_t smem = smem1 > smem2 ? smem1 : smem2;
    if (smem > 200 * 1024) { b2_set_error("sampler: vocabulary %lld too large", (long long)n); return B2_EINVAL; }
    float *cv = (float *)d_scratch;
    int *ci = (int *)((char *)d_scratch + (uint64_t)SAMPLE_BLOCKS * k * 4);
    int *counter = (int *)((char *)d_scratch + (uint64_t)SAMPLE_BLOCKS * k * 8);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_sample_topk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { b2_set_error("sampler: %s", cudaGetErrorString(e)); return B2_ECUDA; }
    }
    k_sample_topk<<<SAMPLE_BLOCKS, 256, smem, st>>>(d_logits, (int)n, k, temperature, top_p, d_u, u_imm, d_out, cv, ci, counter, d_ctx, d_ctx_len,
                                                    ctx_len_imm, penalty, last_n);
    return B2_OK;
}
