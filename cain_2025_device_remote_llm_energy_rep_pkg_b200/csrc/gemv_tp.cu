// gemv_tp.cu — the streaming GEMV with the tensor-parallel exchange fused into it (llama3.1:8b
// decode over 2/4/8 GPUs).  No reference file: Ollama runs one GPU per model here; the call site stays
// experiment/RunnerConfig.py:128-131.
//
// A row-parallel product (O projection, down projection: every rank holds a slice of K) ends in an all-reduce.  It is not a
// kernel of its own and not a library call: the producing GEMV stores each partial row as an 8-byte (value, round tag) pair
// straight into every rank's exchange block (peer memory over NVLink; OUT_TPPART), and the GEMV that consumes the sum reads the
// pairs of this GPU's block until they carry the tag and adds them in rank order while it normalises (IN_TPNORM).  One NVLink
// crossing, no fence, no flag message.  Loads only ever read this GPU's memory; only stores cross the link.
#include "gemv_launch.cuh"

extern int gemv_grid(int n_tiles, int tile_align);
struct TpArgmaxPeers { uint4 *p[8]; };

cudaError_t launch_gemv_tp(const GemvTp &tp, int qt, int in_mode, int out_mode, const void *d_packed, int64_t rows, int64_t cols,
                           const float *norm_w, float eps, int plus_one, const void *act_in, float *y, const float *bias, int act,
                           float *pmax, int *pidx, int *grid_out, cudaStream_t st, bool pdl, int row_off) {
    if (tp.n < 2 || tp.n > 8) return cudaErrorInvalidValue;
    GemvArgsTp a;
    size_t smem = 0;
    cudaError_t em = gemv_make_args(a, in_mode, d_packed, rows, cols, nullptr, norm_w, eps, plus_one, act_in, y, bias, nullptr, act, pmax, pidx,
                                    nullptr, 0, nullptr, 0, nullptr, &smem, qt);
    if (em != cudaSuccess) return em;
    a.pf_ptr = nullptr; a.pf_bytes = 0;
    a.row_off = row_off;
    a.tp_n = tp.n;
    for (int p = 0; p < 8; ++p) a.tp_dst[p] = p < tp.n ? (uint2 *)tp.dst[p] : nullptr;
    a.tp_res = tp.res; a.tp_slots = (const uint2 *)tp.slots; a.tp_slot_stride = tp.slot_stride; a.tp_fwd = tp.fwd;
    a.tp_round = tp.round; a.tp_rounds = tp.rounds; a.tp_xout = tp.xout; a.tp_err = tp.err;
    if (!tp.fwd) return cudaErrorInvalidValue;
    if (in_mode == IN_TPNORM && (!tp.res || !tp.slots || !tp.xout || !tp.err || (cols & 3))) return cudaErrorInvalidValue;
    if (out_mode == OUT_TPPART)
        for (int p = 0; p < tp.n; ++p)
            if (!tp.dst[p]) return cudaErrorInvalidValue;
    const int grid = gemv_grid(a.n_tiles, 1);
    if (a.k_outer && (a.n_tiles + grid - 1) / grid > GEMV_MAXT) return cudaErrorInvalidValue;   // one launch per product
    if (grid_out) *grid_out = grid;
#define CASE(I, O) if (in_mode == I && out_mode == O) return launch_one<I, O>(a, grid, smem, st, pdl)
    CASE(IN_TPNORM, OUT_BIAS); CASE(IN_TPNORM, OUT_ACTMUL); CASE(IN_TPNORM, OUT_LOGITS);
    CASE(IN_ACT, OUT_TPPART); CASE(IN_F32, OUT_TPPART);
#undef CASE
    return cudaErrorInvalidValue;
}

// Closes a forward pass: counts it (the next pass's tags differ) and, when the pass did not run the LM head, first sees the last
// round complete from every rank, so that no rank starts the next pass while another still reads this round's slots.
__global__ void __launch_bounds__(256) k_tp_close(const uint2 *slots, int slot_stride, int n, int n_vec, unsigned long long *fwd, int rounds, int round,
                                                  int wait, unsigned int *err) {
    if (wait) {
        const uint32_t tag = (uint32_t)(*fwd * (unsigned long long)rounds + (unsigned long long)round + 1ull);
        const long long t_begin = global_timer_ns();
        for (int i = threadIdx.x; i < n * (n_vec >> 1); i += blockDim.x) {
            const int p = i / (n_vec >> 1), j = i - p * (n_vec >> 1);
            const uint4 *sp = (const uint4 *)(slots + (size_t)p * slot_stride) + j;
            uint32_t spins = 0;
            for (;;) {
                const uint4 v = ld_volatile_v4(sp);
                if (v.y == tag && v.w == tag) break;
                if ((++spins & 255u) == 0u && (__ldcg(err) != 0u || global_timer_ns() - t_begin > TP_TIMEOUT_NS)) { atomicCAS(err, 0u, 1u + (unsigned)p); break; }
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) *fwd += 1;
}
void launch_tp_close(const void *slots, int slot_stride, int n, int n_vec, unsigned long long *fwd, int rounds, int round, bool wait, unsigned int *err,
                     cudaStream_t st) {
    k_tp_close<<<1, 256, 0, st>>>((const uint2 *)slots, slot_stride, n, n_vec, fwd, rounds, round, wait ? 1 : 0, err);
}

// Greedy sampling on a tensor-parallel group.  Every rank reduces the arg-max partials its LM-head GEMV left (over the whole
// vocabulary when the head ran replicated, over the rank's row window with global indices when it ran vocabulary-parallel),
// stores (value, index, tag, tag) as ONE 16-byte word into slot [this rank] of every rank's block, reads the tp_n slots of its own
// block until they carry the tag, and takes the maximum with the lowest index on ties: the same token on every rank.
__global__ void __launch_bounds__(32) k_tp_argmax(const float *__restrict__ pv, const int *__restrict__ pi, int n, TpArgmaxPeers peers, int tp_n, int rank,
                                                  const uint4 *slots, const unsigned long long *fwd, unsigned int *err, int *__restrict__ out) {
    float v = -INFINITY;
    int idx = 0x7fffffff;
    for (int b = threadIdx.x; b < n; b += 32) arg_better(v, idx, pv[b], pi[b]);
    warp_argmax(v, idx);
    const uint32_t tag = (uint32_t)(__ldcg(fwd) + 1ull) | 0x80000000u;   // never 0; changes with every forward pass
    if ((int)threadIdx.x < tp_n) {
        uint4 *dst = threadIdx.x == 0 ? peers.p[0] : threadIdx.x == 1 ? peers.p[1] : threadIdx.x == 2 ? peers.p[2] : threadIdx.x == 3 ? peers.p[3]
                   : threadIdx.x == 4 ? peers.p[4] : threadIdx.x == 5 ? peers.p[5] : threadIdx.x == 6 ? peers.p[6] : peers.p[7];
        asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + rank), "r"(__float_as_uint(v)), "r"((uint32_t)idx), "r"(tag), "r"(tag) : "memory");
        const long long t_begin = global_timer_ns();
        uint4 w;
        uint32_t spins = 0;
        for (;;) {
            w = ld_volatile_v4(slots + threadIdx.x);
            if (w.z == tag && w.w == tag) break;
            if ((++spins & 255u) == 0u && (__ldcg(err) != 0u || global_timer_ns() - t_begin > TP_TIMEOUT_NS)) { atomicCAS(err, 0u, 1u + threadIdx.x); break; }
        }
        v = __uint_as_float(w.x);
        idx = (int)w.y;
    } else {
        v = -INFINITY;
        idx = 0x7fffffff;
    }
    warp_argmax(v, idx);
    if (threadIdx.x == 0) *out = idx;
}
void launch_tp_argmax(const float *pmax, const int *pidx, int n, void *const *peer_slots, int tp_n, int rank, const void *local_slots,
                      const unsigned long long *fwd, unsigned int *err, int32_t *d_out, cudaStream_t st) {
    TpArgmaxPeers peers;
    for (int p = 0; p < 8; ++p) peers.p[p] = p < tp_n ? (uint4 *)peer_slots[p] : nullptr;
    k_tp_argmax<<<1, 32, 0, st>>>(pmax, pidx, n, peers, tp_n, rank, (const uint4 *)local_slots, fwd, err, d_out);
}

// ------------------------------------------------------------------------------------------ batched prompt evaluation
// The row-parallel prompt GEMMs (O projection, down projection) of a tensor-parallel rank end in the same exchange as the decode
// GEMVs, for a whole batch of tokens: k_pf_tp_scatter adds the GEMM's split-K partials in their fixed order and stores every
// (token, row) sum as a tagged pair into slot [this rank] of EVERY rank's batch buffer (plain stores over NVLink); k_pf_tp_gather
// reads the pairs of this GPU's buffer until they carry the round's tag and forms x = residual + slot_0 + ... + slot_{n-1} in rank
// order, bit-identical on every rank.  Loads only ever read this GPU's memory.
struct PfTpDst { uint2 *p[8]; };
__global__ void __launch_bounds__(256) k_pf_tp_scatter(const float *__restrict__ part, int ksplit, int T, int M, int rows, PfTpDst dst, int tp_n,
                                                       const unsigned long long *fwd, int rounds, int round) {
    const int row = blockIdx.x * 256 + threadIdx.x, m = blockIdx.y;
    if (row >= rows || m >= M) return;
    float v = 0.f;
    for (int s = 0; s < ksplit; ++s) v += part[((int64_t)s * T + m) * rows + row];
    const uint32_t tag = (uint32_t)(__ldcg(fwd) * (unsigned long long)rounds + (unsigned long long)round + 1ull);
    const size_t i = (size_t)m * rows + row;
#pragma unroll
    for (int p = 0; p < 8; ++p)
        if (p < tp_n) st_volatile_v2(dst.p[p] + i, __float_as_uint(v), tag);
}
__global__ void __launch_bounds__(256) k_pf_tp_gather(const uint2 *slots, size_t slot_stride, int tp_n, int M, int rows, const float *res, float *y,
                                                      int64_t ld, const unsigned long long *fwd, int rounds, int round, unsigned int *err) {
    const int row = blockIdx.x * 256 + threadIdx.x, m = blockIdx.y;
    if (row >= rows || m >= M) return;
    const uint32_t tag = (uint32_t)(__ldcg(fwd) * (unsigned long long)rounds + (unsigned long long)round + 1ull);
    const size_t i = (size_t)m * rows + row;
    float v = res[(int64_t)m * ld + row];
    const long long t_begin = global_timer_ns();
    for (int r = 0; r < tp_n; ++r) {
        const uint2 *sp = slots + (size_t)r * slot_stride + i;
        uint2 pr;
        uint32_t spins = 0;
        for (;;) {
            asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(pr.x), "=r"(pr.y) : "l"(sp) : "memory");
            if (pr.y == tag) break;
            if ((++spins & 255u) == 0u && (__ldcg(err) != 0u || global_timer_ns() - t_begin > TP_TIMEOUT_NS)) { atomicCAS(err, 0u, 1u + (unsigned)r); break; }
        }
        v = __fadd_rn(v, __uint_as_float(pr.x));
    }
    y[(int64_t)m * ld + row] = v;
}
void launch_pf_tp_exchange(const float *part, int ksplit, int T, int M, int rows, void *const *peer_slots, const void *local_slots, size_t slot_stride,
                           int tp_n, const float *res, float *y, int64_t ld, const unsigned long long *fwd, int rounds, int round, unsigned int *err,
                           cudaStream_t st) {
    PfTpDst dst;
    for (int p = 0; p < 8; ++p) dst.p[p] = p < tp_n ? (uint2 *)peer_slots[p] : nullptr;
    const dim3 grid((unsigned)((rows + 255) / 256), M);
    k_pf_tp_scatter<<<grid, 256, 0, st>>>(part, ksplit, T, M, rows, dst, tp_n, fwd, rounds, round);
    k_pf_tp_gather<<<grid, 256, 0, st>>>((const uint2 *)local_slots, slot_stride, tp_n, M, rows, res, y, ld, fwd, rounds, round, err);
}
