// ops.cu — op-level C ABI entry points (device pointers, default stream).  They exist so the parity
// tests, bench.py and ncu can drive each stage of the decode path in isolation; the engine (engine.cu)
// calls the same launchers on its own stream.
#include "kernels.cuh"
#include "prefill.cuh"
#include "layout.cuh"
#include "sampler.cuh"

extern "C" uint64_t b2_packed_bytes(int type, int64_t rows, int64_t cols) {
    if (type == B2_T_Q4_K && cols % 256 == 0) return packed_q4_k_bytes(rows, cols);
    if (type == B2_T_Q6_K && cols % 256 == 0) return packed_q6_k_bytes(rows, cols);
    if (type != B2_T_Q4_0 || cols % QK) return 0;
    return packed_q4_0_bytes(rows, cols);
}
extern "C" uint64_t b2_act_bytes(int64_t n) { return a3_bytes(n); }

extern "C" int b2_op_pack_weights(int type, int64_t rows, int64_t cols, const void *src, const void *src2, void *d_packed) {
    ApiScope _scope("b2_op_pack_weights");
    B2_CHECK(type == B2_T_Q4_0 || type == B2_T_Q4_K || type == B2_T_Q6_K, B2_EINVAL, "b2_op_pack_weights: type %d not supported", type);
    const bool kq = type != B2_T_Q4_0;
    B2_CHECK(src && d_packed && rows > 0 && cols > 0 && cols % (kq ? 256 : QK) == 0, B2_EINVAL, "b2_op_pack_weights: bad arguments");
    uint64_t bytes = type == B2_T_Q6_K ? (uint64_t)rows * (cols / 256) * 210 : kq ? (uint64_t)rows * (cols / 256) * 144 : (uint64_t)rows * (cols / QK) * 18;
    auto pack = [&](void *stg, int64_t row0, int64_t stride) {
        if (type == B2_T_Q6_K) launch_pack_q6_k(stg, rows, cols, d_packed, row0, stride, 0);
        else if (kq) launch_pack_q4_k(stg, rows, cols, d_packed, row0, stride, 0);
        else launch_pack_q4_0(stg, rows, cols, d_packed, row0, stride, 0);
    };
    int64_t total = src2 ? 2 * rows : rows, stride = src2 ? 2 : 1;
    void *stg = nullptr;
    B2_CUDA(cudaMalloc(&stg, bytes));
    cudaError_t e = cudaMemset(d_packed, 0, b2_packed_bytes(type, total, cols));
    if (e == cudaSuccess) e = cudaMemcpy(stg, src, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        pack(stg, 0, stride);
        e = cudaDeviceSynchronize();
    }
    if (e == cudaSuccess && src2) {
        e = cudaMemcpy(stg, src2, bytes, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) {
            pack(stg, 1, stride);
            e = cudaDeviceSynchronize();
        }
    }
    cudaFree(stg);
    B2_CUDA(e);
    return B2_OK;
}

extern "C" int b2_op_quant_act(const float *d_x, int64_t n, void *d_act) {
    ApiScope _scope("b2_op_quant_act");
    B2_CHECK(d_x && d_act && n >= 0 && n % QK == 0, B2_EINVAL, "b2_op_quant_act: n must be a multiple of 32");
    if (n == 0) return B2_OK;
    B2_CUDA(cudaMemsetAsync(d_act, 0, a3_bytes(n), 0));
    launch_quant_act(d_x, n, d_act, 0);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
extern "C" int b2_op_rmsnorm_quant(const float *d_x, const float *d_w, int64_t n, float eps, int plus_one, void *d_act, float *d_y_opt) {
    ApiScope _scope("b2_op_rmsnorm_quant");
    B2_CHECK(d_x && d_w && d_act && n > 0 && n % QK == 0 && n <= 16384, B2_EINVAL, "b2_op_rmsnorm_quant: n must be a multiple of 32 in (0,16384]");
    B2_CUDA(cudaMemsetAsync(d_act, 0, a3_bytes(n), 0));
    launch_rmsnorm_quant(d_x, d_w, n, eps, plus_one, d_act, d_y_opt, 0);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
extern "C" int b2_op_gemv(int type, const void *d_packed, int64_t rows, int64_t cols, const void *d_act, const float *d_bias_or_res,
                          int mode, int act, float *d_y) {
    ApiScope _scope("b2_op_gemv");
    B2_CHECK(type == B2_T_Q4_0 || type == B2_T_Q4_K || type == B2_T_Q6_K, B2_EINVAL, "b2_op_gemv: type %d not supported", type);
    const int qt = qt_of(type);
    B2_CHECK(d_packed && d_act && d_y && rows >= 0 && cols > 0 && cols % QK == 0, B2_EINVAL, "b2_op_gemv: bad arguments");
    B2_CHECK(mode >= 0 && mode <= 2, B2_EINVAL, "b2_op_gemv: mode %d", mode);
    B2_CHECK(mode != 1 || d_bias_or_res, B2_EINVAL, "b2_op_gemv: mode 1 needs the residual vector");
    B2_CHECK(mode != 2 || rows % 2 == 0, B2_EINVAL, "b2_op_gemv: mode 2 needs an even row count (gate/up pairs)");
    if (rows == 0) return B2_OK;
    cudaError_t e = launch_gemv_qt(qt, GEMV_IN_ACT, mode, d_packed, rows, cols, nullptr, nullptr, 0.f, 0, d_act, d_y, d_bias_or_res, nullptr, act,
                                nullptr, nullptr, nullptr, nullptr, 0, nullptr, 0, nullptr, 0, false);
    B2_CUDA(e);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
// y = W * grid(x) (+bias / +res) with the fp32 input vector put on the activation grid inside the kernel (the engine's down
// projection consumes act(gate)*up this way).  mode 0 or 1 as b2_op_gemv.
extern "C" int b2_op_gemv_f32(int type, const void *d_packed, int64_t rows, int64_t cols, const float *d_x, const float *d_bias_or_res,
                              int mode, float *d_y) {
    ApiScope _scope("b2_op_gemv_f32");
    B2_CHECK(type == B2_T_Q4_0 || type == B2_T_Q4_K || type == B2_T_Q6_K, B2_EINVAL, "b2_op_gemv_f32: type %d not supported", type);
    const int qt = qt_of(type);
    B2_CHECK(d_packed && d_x && d_y && rows > 0 && cols > 0 && cols % QK == 0, B2_EINVAL, "b2_op_gemv_f32: bad arguments");
    B2_CHECK(mode == 0 || (mode == 1 && d_bias_or_res), B2_EINVAL, "b2_op_gemv_f32: mode must be 0, or 1 with a residual vector");
    cudaError_t e = launch_gemv_qt(qt, GEMV_IN_F32, mode, d_packed, rows, cols, nullptr, nullptr, 0.f, 0, d_x, d_y, d_bias_or_res, nullptr, 0,
                                nullptr, nullptr, nullptr, nullptr, 0, nullptr, 0, nullptr, 0, false);
    B2_CHECK(e != cudaErrorInvalidValue, B2_EINVAL, "b2_op_gemv_f32: cols %lld too long for the shared-memory staging", (long long)cols);
    B2_CUDA(e);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
// Fused form used by the engine: RMSNorm + activation grid + GEMV (+ epilogue).  mode 0..2 as b2_op_gemv; mode 3
// writes logits to d_y and the index of their maximum to d_argmax.  d_act_out (mode 2, optional) receives the
// activation-grid image of the result (b2_act_bytes(rows/2) bytes).
extern "C" int b2_op_norm_gemv(int type, const void *d_packed, int64_t rows, int64_t cols, const float *d_x, const float *d_norm_w, float eps,
                               int plus_one, const float *d_bias_or_res, int mode, int act, float *d_y, void *d_act_out, int32_t *d_argmax) {
    ApiScope _scope("b2_op_norm_gemv");
    B2_CHECK(type == B2_T_Q4_0 || type == B2_T_Q4_K || type == B2_T_Q6_K, B2_EINVAL, "b2_op_norm_gemv: type %d not supported", type);
    const int qt = qt_of(type);
    B2_CHECK(d_packed && d_x && d_norm_w && rows > 0 && cols > 0 && cols % QK == 0 && cols <= 8192, B2_EINVAL, "b2_op_norm_gemv: bad arguments (cols must be a multiple of 32, at most 8192)");
    B2_CHECK(mode >= 0 && mode <= 3, B2_EINVAL, "b2_op_norm_gemv: mode %d", mode);
    B2_CHECK(d_y, B2_EINVAL, "b2_op_norm_gemv: d_y is NULL");
    B2_CHECK(mode != 1 || d_bias_or_res, B2_EINVAL, "b2_op_norm_gemv: mode 1 needs the residual vector");
    B2_CHECK(mode != 2 || rows % 64 == 0, B2_EINVAL, "b2_op_norm_gemv: mode 2 needs rows %% 64 == 0 (32 gate/up pairs per activation block)");
    B2_CHECK(mode != 3 || d_argmax, B2_EINVAL, "b2_op_norm_gemv: mode 3 needs d_argmax");
    void *tmp = nullptr;
    B2_CUDA(cudaMalloc(&tmp, 8192 + 16));
    B2_CUDA(cudaMemset(tmp, 0, 8192 + 16));
    float *pmax = (float *)tmp;
    int *pidx = (int *)((char *)tmp + 4096);
    float *xsq = (float *)((char *)tmp + 8192);
    launch_sumsq(d_x, cols, xsq, 0);   // the engine gets this sum from the kernel that produced x
    int grid = 0;
    cudaError_t e = launch_gemv_qt(qt, GEMV_IN_NORM, mode, d_packed, rows, cols, d_x, d_norm_w, eps, plus_one, nullptr, d_y, d_bias_or_res, nullptr, act,
                                mode == 3 ? pmax : nullptr, mode == 3 ? pidx : nullptr, &grid, nullptr, 0, xsq, 1, nullptr, 0, false);
    if (e == cudaSuccess && mode == 3) launch_argmax_partials(pmax, pidx, grid, d_argmax, 0);
    if (e == cudaSuccess && mode == 2 && d_act_out) {   // the engine's next GEMV puts h on the grid itself; offered here for callers that want the image
        e = cudaMemsetAsync(d_act_out, 0, a3_bytes(rows / 2), 0);
        if (e == cudaSuccess) launch_quant_act(d_y, rows / 2, d_act_out, 0);
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaFree(tmp);
    B2_CUDA(e);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
extern "C" int b2_op_rope_kv(float *d_qkv, int nh, int nkv, int hd, int pos, const float *d_cos, const float *d_sin, void *d_kcache,
                             void *d_vcache, const int32_t *d_page_table, int page_tokens) {
    ApiScope _scope("b2_op_rope_kv");
    B2_CHECK(page_tokens == PAGE_TOKENS, B2_EINVAL, "b2_op_rope_kv: page_tokens must be %d", PAGE_TOKENS);
    B2_CHECK(d_qkv && d_cos && d_sin && d_kcache && d_vcache && d_page_table && pos >= 0, B2_EINVAL, "b2_op_rope_kv: bad arguments");
    B2_CHECK(hd % 2 == 0 && nh > 0 && nkv > 0, B2_EINVAL, "b2_op_rope_kv: bad head geometry");
    launch_rope_kv(d_qkv, nh, nkv, hd, nullptr, pos, d_cos, d_sin, (__half *)d_kcache, (__half *)d_vcache, d_page_table, 0);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
extern "C" int b2_op_attn_decode(const float *d_q, int nh, int nkv, int hd, int n_ctx, const void *d_kcache, const void *d_vcache,
                                 const int32_t *d_page_table, int page_tokens, float *d_out) {
    ApiScope _scope("b2_op_attn_decode");
    B2_CHECK(page_tokens == PAGE_TOKENS, B2_EINVAL, "b2_op_attn_decode: page_tokens must be %d", PAGE_TOKENS);
    B2_CHECK(d_q && d_kcache && d_vcache && d_page_table && d_out, B2_EINVAL, "b2_op_attn_decode: NULL pointer");
    B2_CHECK(n_ctx > 0 && n_ctx <= 65536, B2_EINVAL, "b2_op_attn_decode: n_ctx %d out of range", n_ctx);
    B2_CHECK(hd % 32 == 0 && hd >= 32 && hd <= 256 && nh > 0 && nkv > 0 && nh % nkv == 0, B2_EINVAL, "b2_op_attn_decode: bad head geometry");
    void *scratch = nullptr;
    uint64_t sb = attn_scratch_bytes(nh, hd);
    B2_CUDA(cudaMalloc(&scratch, sb));
    cudaError_t e = cudaMemset(scratch, 0, sb);
    if (e == cudaSuccess) {
        launch_attn_decode(d_q, nh, nkv, hd, nullptr, n_ctx, (const __half *)d_kcache, (const __half *)d_vcache, d_page_table, d_out, scratch, n_ctx, 0);
        e = cudaDeviceSynchronize();
    }
    cudaFree(scratch);
    B2_CUDA(e);
    return B2_OK;
}
extern "C" int b2_op_attn_fused(const float *d_qkv, int nh, int nkv, int hd, int pos, const float *d_cos, const float *d_sin, void *d_kcache,
                                void *d_vcache, const int32_t *d_page_table, int page_tokens, int max_ctx, float *d_out, void *d_act_out) {
    ApiScope _scope("b2_op_attn_fused");
    B2_CHECK(page_tokens == PAGE_TOKENS, B2_EINVAL, "b2_op_attn_fused: page_tokens must be %d", PAGE_TOKENS);
    B2_CHECK(d_qkv && d_cos && d_sin && d_kcache && d_vcache && d_page_table && (d_out || d_act_out), B2_EINVAL, "b2_op_attn_fused: NULL pointer");
    B2_CHECK(pos >= 0 && pos < max_ctx && max_ctx <= 65536, B2_EINVAL, "b2_op_attn_fused: pos %d / max_ctx %d out of range", pos, max_ctx);
    B2_CHECK(hd % 32 == 0 && hd >= 32 && hd <= 256 && nh > 0 && nkv > 0 && nh % nkv == 0, B2_EINVAL, "b2_op_attn_fused: bad head geometry");
    B2_CUDA(launch_attn_fused(d_qkv, nh, nkv, hd, nullptr, pos, d_cos, d_sin, (__half *)d_kcache, (__half *)d_vcache, d_page_table, d_out,
                              d_act_out, max_ctx, 0, false));
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
extern "C" int b2_op_argmax(const float *d_logits, int64_t n, int32_t *d_out_index) {
    ApiScope _scope("b2_op_argmax");
    // the kernel strides an int index by its grid (a few thousand threads): stay that far below 2^31
    B2_CHECK(d_logits && d_out_index && n > 0 && n <= (1ll << 31) - (1ll << 20), B2_EINVAL, "b2_op_argmax: bad arguments (n must be in [1, 2^31 - 2^20])");
    void *scratch = nullptr;
    B2_CUDA(cudaMalloc(&scratch, argmax_scratch_bytes()));
    cudaError_t e = cudaMemset(scratch, 0, argmax_scratch_bytes());
    if (e == cudaSuccess) {
        launch_argmax(d_logits, n, d_out_index, scratch, 0);
        e = cudaDeviceSynchronize();
    }
    cudaFree(scratch);
    B2_CUDA(e);
    return B2_OK;
}
extern "C" int b2_op_sample_topk(const float *d_logits, int64_t n, int top_k, float temperature, float top_p, float u, int32_t *d_out_index) {
    ApiScope _scope("b2_op_sample_topk");
    B2_CHECK(d_logits && d_out_index && n > 0 && n < (1ll << 31), B2_EINVAL, "b2_op_sample_topk: bad arguments");
    B2_CHECK(top_k >= 1 && top_k <= SAMPLE_MAX_K, B2_EINVAL, "b2_op_sample_topk: top_k %d out of range [1,%d]", top_k, SAMPLE_MAX_K);
    void *scratch = nullptr;
    uint64_t sb = sample_scratch_bytes(top_k);
    B2_CUDA(cudaMalloc(&scratch, sb));
    cudaError_t e = cudaMemset(scratch, 0, sb);
    int rc = B2_OK;
    if (e == cudaSuccess) {
        rc = launch_sample_topk(d_logits, n, top_k, temperature, top_p, nullptr, u, d_out_index, scratch, 0);
        e = cudaDeviceSynchronize();
    }
    cudaFree(scratch);
    if (rc) return rc;
    B2_CUDA(e);
    return B2_OK;
}

// The same with the repetition penalty of Ollama's options.repeat_penalty / repeat_last_n: d_ctx holds the ctx_len tokens of the
// context so far; every distinct token among its last `repeat_last_n` has its logit divided by the penalty when positive and
// multiplied by it otherwise before the candidates are selected.
extern "C" int b2_op_sample_topk_penalty(const float *d_logits, int64_t n, int top_k, float temperature, float top_p, float u, const int32_t *d_ctx,
                                         int ctx_len, float repeat_penalty, int repeat_last_n, int32_t *d_out_index) {
    ApiScope _scope("b2_op_sample_topk_penalty");
    B2_CHECK(d_logits && d_out_index && n > 0 && n < (1ll << 31), B2_EINVAL, "b2_op_sample_topk_penalty: bad arguments");
    B2_CHECK(top_k >= 1 && top_k <= SAMPLE_MAX_K, B2_EINVAL, "b2_op_sample_topk_penalty: top_k %d out of range [1,%d]", top_k, SAMPLE_MAX_K);
    B2_CHECK(ctx_len >= 0 && (d_ctx || ctx_len == 0) && repeat_penalty >= 0.f && repeat_last_n >= 0, B2_EINVAL, "b2_op_sample_topk_penalty: bad context / penalty");
    void *scratch = nullptr;
    uint64_t sb = sample_scratch_bytes(top_k);
    B2_CUDA(cudaMalloc(&scratch, sb));
    cudaError_t e = cudaMemset(scratch, 0, sb);
    int rc = B2_OK;
    if (e == cudaSuccess) {
        rc = launch_sample_topk(d_logits, n, top_k, temperature, top_p, nullptr, u, d_out_index, scratch, 0, d_ctx, nullptr, ctx_len, repeat_penalty, repeat_last_n);
        e = cudaDeviceSynchronize();
    }
    cudaFree(scratch);
    if (rc) return rc;
    B2_CUDA(e);
    return B2_OK;
}

// Batched form of b2_op_norm_gemv / b2_op_gemv_f32 for M tokens (the prompt-evaluation GEMM on tcgen05 tensor cores, prefill.cu):
// y[m] = W * grid(RMSNorm(x[m]))  (d_norm_w NULL: no norm) with mode 0 (+bias[rows], may be NULL), 1 (+res[m][rows]) or
// 2 (act(gate)*up, y has rows/2 columns).  x is [M][cols], y [M][rows or rows/2], both dense row-major.
extern "C" int b2_op_prefill_gemm(int type, const void *d_packed, int64_t rows, int64_t cols, const float *d_x, int M, const float *d_norm_w,
                                  float eps, int plus_one, const float *d_bias_or_res, int mode, int act, float *d_y) {
    ApiScope _scope("b2_op_prefill_gemm");
    B2_CHECK(type == B2_T_Q4_0 || type == B2_T_Q4_K || type == B2_T_Q6_K, B2_EINVAL, "b2_op_prefill_gemm: type %d not supported", type);
    B2_CHECK(type == B2_T_Q4_0 || cols % 256 == 0, B2_EINVAL, "b2_op_prefill_gemm: K-quants need cols %% 256 == 0");
    B2_CHECK(d_packed && d_x && d_y && rows > 0 && cols > 0 && cols % QK == 0 && M > 0, B2_EINVAL, "b2_op_prefill_gemm: bad arguments");
    B2_CHECK(mode >= 0 && mode <= 2, B2_EINVAL, "b2_op_prefill_gemm: mode %d", mode);
    B2_CHECK(mode != 1 || d_bias_or_res, B2_EINVAL, "b2_op_prefill_gemm: mode 1 needs the residual matrix");
    B2_CHECK(mode != 2 || rows % 2 == 0, B2_EINVAL, "b2_op_prefill_gemm: mode 2 needs an even number of rows");
    int dev = 0, sms = 0;
    B2_CUDA(cudaGetDevice(&dev));
    B2_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    void *pb = nullptr;
    float *part = nullptr;
    const uint64_t part_floats = (uint64_t)8 * PF_MAX_TOKENS * rows;
    B2_CUDA(cudaMalloc(&pb, pb_bytes(PF_MAX_TOKENS, cols)));
    if (cudaMalloc((void **)&part, part_floats * 4) != cudaSuccess) { cudaFree(pb); cudaGetLastError(); b2_set_error("b2_op_prefill_gemm: out of memory"); return B2_ENOMEM; }
    const int64_t ldy = mode == 2 ? rows / 2 : rows;
    cudaError_t e = cudaSuccess;
    for (int m0 = 0; m0 < M && e == cudaSuccess; m0 += PF_MAX_TOKENS) {
        const int mb = M - m0 < PF_MAX_TOKENS ? M - m0 : PF_MAX_TOKENS, T = pf_batch(mb);
        e = launch_prefill_quant(d_norm_w ? PF_IN_NORM : PF_IN_F32, d_x + (int64_t)m0 * cols, cols, mb, T, cols, d_norm_w, eps, plus_one, pb, 0);
        if (e == cudaSuccess)
            e = launch_prefill_gemm_qt(qt_of(type), mode, d_packed, rows, cols, pb, mb, T, d_y + (int64_t)m0 * ldy, ldy,
                                    mode == 1 ? d_bias_or_res + (int64_t)m0 * ldy : d_bias_or_res, act, sms, part, part_floats, 0);
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaFree(pb);
    cudaFree(part);
    B2_CUDA(e);
    return B2_OK;
}
