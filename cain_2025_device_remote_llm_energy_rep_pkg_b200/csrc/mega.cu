// mega.cu — the persistent token kernel: the whole decode loop of Ollama's /api/generate for
// any number of consecutive tokens in ONE cooperative launch of one 512-thread CTA per SM.
//
// Why: with one kernel per stage (gemv.cu, attn.cu under a CUDA graph with programmatic dependent launch) a layer of
// llama3.1:8b costs ~53 us of which ~22 us stream weights; the rest is per-kernel overhead (launch hand-over ~2 us, CTA
// start-up and tail ~2 us, four GEMVs + attention per layer).  Here a stage boundary is a grid-wide barrier (~1.2 us): the
// CTAs stay resident, every CTA requests the first weight pieces of the next stage BEFORE it waits at the barrier (weights
// never depend on the previous stage), and only the small activation vector is fetched after it.
//
// Stages per token (each a "phase" between two grid barriers):
//   token begin: arg-max over the per-CTA partials of the previous LM head (or the next prompt token), optional copy of the
//                logits to the log, embedding row -> x and partial sums of x^2          (every CTA redundantly / sliced)
//   per layer:   QKV GEMV (RMSNorm fused) | attention (four 128-thread groups per CTA, one (head, split) item each)
//                | O-projection GEMV (+residual) | gate/up GEMV (RMSNorm, act*mul) | down GEMV (+residual)
//   LM head GEMV (RMSNorm, logits + arg-max partials)
// The arithmetic is the code of gemv_dev.cuh / attn_dev.cuh, shared with the stand-alone kernels.
#include "mega.cuh"

__constant__ MegaPlan c_plan;

template <bool LOG>
__device__ __forceinline__ void copy_logits(const float *src, float *dst, int n) {
    const int n4 = n >> 2;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x)
        ((float4 *)dst)[i] = __ldcg((const float4 *)src + i);
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) dst[n4 * 4 + threadIdx.x] = __ldcg(src + n4 * 4 + threadIdx.x);
}

enum { ST_QKV = 0, ST_O, ST_GATEUP, ST_DOWN, ST_HEAD };
// A stage of the token loop.  Its argument block sits at a fixed constant-bank address: the fields are instruction operands.
template <int IN, int OUT, int ST>
__device__ __forceinline__ void gemv_stage(const uint8_t *w, const float *norm_w, const float *bias, uint8_t *smem, uint32_t &parity) {
    const GemvArgs &a = ST == ST_QKV ? c_plan.qkv : ST == ST_O ? c_plan.o : ST == ST_GATEUP ? c_plan.gateup : ST == ST_DOWN ? c_plan.down : c_plan.head;
    const GemvPtrs pv = {ST == ST_HEAD ? a.w : w, (ST == ST_QKV || ST == ST_GATEUP) ? norm_w : a.norm_w, ST == ST_QKV ? bias : a.bias_or_res};
    gemv_phase<IN, OUT, true>(a, pv, smem, parity);
}

__global__ void __launch_bounds__(GEMV_THREADS, 1) k_token_loop(const MegaRun r) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ int s_tok;
    const MegaPlan &P = c_plan;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float *scratch = (float *)smem;              // the scratch floats of gemv_phase
    uint64_t *bar = (uint64_t *)(smem + 256);    // its input-copy mbarrier
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
#ifdef MEGA_PROF
    if (tid == 0) s_prof_seq = 0;
#endif
    __syncthreads();
    uint32_t parity = 0;
    for (int i = 0; i < r.n_tokens; ++i) {
        const int pos = r.pos0 + i;
        if (i > 0) grid_barrier();   // the LM head (or the last down-projection) of the previous token is complete
        // ---------------- token begin
        prof_stamp(0);
        int cur;
        if (r.mode == MEGA_DECODE) {
            if (warp == 0) {
                float bv = -INFINITY;
                int bi = 0x7fffffff;
                for (int k = lane; k < P.n_part; k += 32) {
                    const float v = __ldcg(P.pmax + k);
                    const int ix = __ldcg(P.pidx + k);
                    if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                    if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
                }
                if (lane == 0) s_tok = bi;
            }
            __syncthreads();
            cur = s_tok;
            if (blockIdx.x == 0 && tid == 0) r.out_tokens[i] = cur;
            if (r.logits_log) copy_logits<true>(P.logits, r.logits_log + (int64_t)i * P.vocab, P.vocab);
        } else {
            cur = __ldg(r.prompt + i);
            if (r.logits_log && i > 0 && i - 1 >= r.head_from) copy_logits<true>(P.logits, r.logits_log + (int64_t)(i - 1) * P.vocab, P.vocab);
        }
        {   // embedding row: 32-element blocks dealt round-robin to the warps of the grid; per-CTA partial of sum(x^2)
            float ss = 0.f;
            const int nblk = P.n_embd / QK;
            for (int b = blockIdx.x * GEMV_CWARPS + warp; b < nblk; b += gridDim.x * GEMV_CWARPS) {
                float v = packed_elem(P.embd_w, cur, (int64_t)b * QK + lane, P.embd_nbp);
                if (P.embd_scale != 1.0f) v = v * P.embd_scale;
                P.x[b * QK + lane] = v;
                ss = fmaf(v, v, ss);
            }
            ss = warp_sum(ss);
            __syncthreads();   // s_tok / scratch: everybody is past the previous use
            if (lane == 0) scratch[warp] = ss;
            __syncthreads();
            if (tid == 0) {
                float tot = 0.f;
#pragma unroll
                for (int w = 0; w < GEMV_CWARPS; ++w) tot += scratch[w];
                P.xsq[blockIdx.x] = tot;
            }
            prof_stamp(5);
            prof_next();
        }
        // ---------------- layers
        for (int l = 0; l < P.n_layers; ++l) {
            const MegaLayer &L = P.layers[l];
            {
                gemv_stage<IN_NORM, OUT_BIAS, ST_QKV>(L.qkv_w, L.attn_norm, L.qkv_bias, smem, parity);
            }
            {
                AttnArgs A = P.attn;
                A.kc = L.kc; A.vc = L.vc;
                const int nsplit = (pos + 1 <= ATTN_SHORT_CTX) ? 1 : ATTN_NSPLIT;
                const int grp = warp >> 2, gt = tid & 127;
                const int item = grp * gridDim.x + blockIdx.x;          // items spread over the SMs first, then over the groups of a CTA
                const bool active = item < A.nh * nsplit;
                const int h = item / nsplit, split = item - h * nsplit;
                const AttnSmem as = attn_carve((float *)(smem + GEMV_FIXED_SMEM(GEMV_RB_MAX)) + grp * attn_group_floats(A.hd, A.pps_max), A.hd, A.pps_max);
                __syncthreads();   // every warp of this CTA has left the QKV phase: its shared memory is free
                prof_stamp(0);
                if (active) attn_pages(A, as, gt, h, split, nsplit, pos);
                prof_stamp(1);
                grid_barrier();    // q, k, v of this token are complete
                prof_stamp(2);
                if (active) { AttnRegs R; attn_main<MERGE_GLOBAL, 0>(A, as, gt, h, split, nsplit, pos, 1 + grp, R); }
                prof_stamp(5);
                prof_next();
            }
            {
                gemv_stage<IN_ACT, OUT_RES, ST_O>(L.o_w, nullptr, nullptr, smem, parity);
            }
            {
                gemv_stage<IN_NORM, OUT_ACTMUL, ST_GATEUP>(L.gateup_w, L.ffn_norm, nullptr, smem, parity);
            }
            {
                gemv_stage<IN_F32, OUT_RES, ST_DOWN>(L.down_w, nullptr, nullptr, smem, parity);
            }
        }
        if (i >= r.head_from) gemv_stage<IN_NORM, OUT_LOGITS, ST_HEAD>(nullptr, nullptr, nullptr, smem, parity);
    }
    if (r.mode == MEGA_PROMPT && r.logits_log && r.n_tokens - 1 >= r.head_from) {
        grid_barrier();
        copy_logits<true>(P.logits, r.logits_log + (int64_t)(r.n_tokens - 1) * P.vocab, P.vocab);
    }
    if (blockIdx.x == 0 && tid == 0 && r.state) {
        r.state->pos = r.pos0 + r.n_tokens;
        r.state->step = r.n_tokens;
    }
}

// ------------------------------------------------------------------------------------------ host side
static const void *g_plan_owner[64] = {};   // per device: whose plan the constant copy holds

void mega_forget(const void *owner) {
    for (auto &o : g_plan_owner)
        if (o == owner) o = nullptr;
}

cudaError_t mega_grid(size_t smem, int *grid_out) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_token_loop, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_token_loop, GEMV_THREADS, smem)) != cudaSuccess) return e;
    if (per_sm < 1 || sms < 1) return cudaErrorLaunchOutOfResources;
    *grid_out = sms;   // one CTA per SM: the streaming loop is written for a CTA that owns its SM
    return cudaSuccess;
}

cudaError_t launch_mega(const MegaPlan *host_plan, const void *owner, const MegaRun &run, int grid, size_t smem, cudaStream_t st) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (g_plan_owner[dev] != owner) {
        // stream-ordered: kernels already queued on `st` keep the old copy until they finish.  (One stream per device is assumed,
        // as in b2_engine; two engines on one device must not interleave sessions without synchronising.)
        const size_t used = offsetof(MegaPlan, layers) + (size_t)host_plan->n_layers * sizeof(MegaLayer);
        if ((e = cudaMemcpyToSymbolAsync(c_plan, host_plan, used, 0, cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
        g_plan_owner[dev] = owner;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GEMV_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, k_token_loop, run);
}

// development aid: stamp buffer of the profiling build (EXTRA=-DMEGA_PROF); -1 in a normal build
extern "C" int b2dbg_mega_profile(long long *d_buf, int cap_stages) {
#ifdef MEGA_PROF
    if (cudaMemcpyToSymbol(g_prof, &d_buf, sizeof d_buf) != cudaSuccess) return -2;
    if (cudaMemcpyToSymbol(g_prof_cap, &cap_stages, sizeof cap_stages) != cudaSuccess) return -2;
    return 0;
#else
    (void)d_buf; (void)cap_stages;
    return -1;
#endif
}
