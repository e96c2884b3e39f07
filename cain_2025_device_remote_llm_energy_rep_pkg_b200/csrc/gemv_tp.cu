// gemv_tp.cu — the streaming GEMV with the tensor-parallel exchange fused into it (llama3.1:8b
// decode over 2/4/8 GPUs).  No reference file: Ollama runs one GPU per model here; the call site stays
// experiment/RunnerConfig.py:128-131.
//
// A row-parallel product (O projection, down projection: every rank holds a slice of K) ends in an all-reduce.  It is not a
// kernel of its own and not a library call: the producing GEMV stores its partial rows straight into every rank's exchange
// buffer (peer memory over NVLink, plain stores + one arrival count per CTA and rank: OUT_TPPART), and the GEMV that consumes
// the sum waits for the counts, copies residual and slots from its OWN memory into shared memory and adds them in rank order
// while it normalises (IN_TPNORM).  Bulk copies only ever read this GPU's memory.
#include "gemv_launch.cuh"

extern int gemv_grid(int n_tiles, int tile_align);

cudaError_t launch_gemv_tp(const GemvTp &tp, int qt, int in_mode, int out_mode, const void *d_packed, int64_t rows, int64_t cols,
                           const float *norm_w, float eps, int plus_one, const void *act_in, float *y, const float *bias, int act,
                           float *pmax, int *pidx, int *grid_out, cudaStream_t st, bool pdl) {
    if (tp.n < 2 || tp.n > 8) return cudaErrorInvalidValue;
    GemvArgsTp a;
    size_t smem = 0;
    cudaError_t em = gemv_make_args(a, in_mode, d_packed, rows, cols, nullptr, norm_w, eps, plus_one, act_in, y, bias, nullptr, act, pmax, pidx,
                                    nullptr, 0, nullptr, 0, nullptr, &smem, qt);
    if (em != cudaSuccess) return em;
    a.pf_ptr = nullptr; a.pf_bytes = 0;
    a.tp_n = tp.n;
    for (int p = 0; p < 8; ++p) { a.tp_dst[p] = p < tp.n ? tp.dst[p] : nullptr; a.tp_sig[p] = p < tp.n ? tp.sig[p] : nullptr; }
    a.tp_res = tp.res; a.tp_slots = tp.slots; a.tp_slot_stride = tp.slot_stride; a.tp_cnt = tp.cnt; a.tp_fwd = tp.fwd;
    a.tp_round = tp.round; a.tp_rounds = tp.rounds; a.tp_per_round = tp.per_round; a.tp_xout = tp.xout; a.tp_err = tp.err;
    if (in_mode == IN_TPNORM) {
        if (!tp.res || !tp.slots || !tp.cnt || !tp.fwd || !tp.xout || !tp.err || tp.per_round <= 0 || (cols & 3)) return cudaErrorInvalidValue;
        smem += (size_t)tp.n * cols * 4;
        if (smem > 227 * 1024 && a.rb > 4) {   // make room for the slots: park fewer finished tiles between reductions
            smem -= (size_t)(a.rb - 4) * 4096;
            a.rb = 4;
        }
        if (smem > 227 * 1024) return cudaErrorInvalidValue;
    }
    if (out_mode == OUT_TPPART)
        for (int p = 0; p < tp.n; ++p)
            if (!tp.dst[p] || !tp.sig[p]) return cudaErrorInvalidValue;
    const int grid = gemv_grid(a.n_tiles, 1);
    if (a.k_outer && (a.n_tiles + grid - 1) / grid > GEMV_MAXT) return cudaErrorInvalidValue;   // one launch per product: the arrival counts assume it
    if (grid_out) *grid_out = grid;
#define CASE(I, O) if (in_mode == I && out_mode == O) return launch_one<I, O>(a, grid, smem, st, pdl)
    CASE(IN_TPNORM, OUT_BIAS); CASE(IN_TPNORM, OUT_ACTMUL); CASE(IN_TPNORM, OUT_LOGITS);
    CASE(IN_ACT, OUT_TPPART); CASE(IN_F32, OUT_TPPART);
#undef CASE
    return cudaErrorInvalidValue;
}

// One thread: waits until the arrival counters of a round are complete (closes a forward pass that does not run the LM head,
// so that no rank starts the next pass while another still reads this round's slots) and counts the finished pass.
__global__ void k_tp_close(const unsigned long long *cnt, int n, unsigned long long *fwd, int rounds, int round, int per_round, int wait,
                           unsigned int *err) {
    if (wait) {
        const unsigned long long target = (unsigned long long)per_round * (*fwd * (unsigned long long)rounds + (unsigned long long)round + 1ull);
        const long long t_begin = global_timer_ns();
        for (int p = 0; p < n && *err == 0u; ++p)
            while (ld_acquire_sys_u64(cnt + p) < target)
                if (global_timer_ns() - t_begin > TP_TIMEOUT_NS) { atomicExch(err, 1u + (unsigned)p); break; }
    }
    *fwd += 1;
}
void launch_tp_close(const unsigned long long *cnt, int n, unsigned long long *fwd, int rounds, int round, int per_round, bool wait,
                     unsigned int *err, cudaStream_t st) {
    k_tp_close<<<1, 1, 0, st>>>(cnt, n, fwd, rounds, round, per_round, wait ? 1 : 0, err);
}
