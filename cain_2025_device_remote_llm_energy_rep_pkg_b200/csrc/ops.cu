// ops.cu — op-level C ABI entry points (device pointers, default stream).  They exist so the parity
// tests, bench.py and ncu can drive each stage of the decode path in isolation; the engine (engine.cu)
// calls the same launchers on its own stream.
#include "kernels.cuh"
#include "sampler.cuh"

extern "C" uint64_t b2_packed_bytes(int type, int64_t rows, int64_t cols) {
    if (type != B2_T_Q4_0 || cols % QK) return 0;
    return q4_0_packed_bytes(rows, cols);
}
extern "C" uint64_t b2_act_bytes(int64_t n) { return act_bytes(n); }

extern "C" int b2_op_pack_weights(int type, int64_t rows, int64_t cols, const void *src, const void *src2, void *d_packed) {
    ApiScope _scope("b2_op_pack_weights");
    B2_CHECK(type == B2_T_Q4_0, B2_EINVAL, "b2_op_pack_weights: type %d not supported", type);
    B2_CHECK(src && d_packed && rows > 0 && cols > 0 && cols % QK == 0, B2_EINVAL, "b2_op_pack_weights: bad arguments");
    uint64_t bytes = (uint64_t)rows * (cols / QK) * 18;
    void *stg = nullptr;
    B2_CUDA(cudaMalloc(&stg, bytes));
    int64_t total = src2 ? 2 * rows : rows, stride = src2 ? 2 : 1;
    cudaError_t e = cudaMemcpy(stg, src, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        launch_pack_q4_0(stg, rows, cols, d_packed, 0, stride, total, 0);
        e = cudaDeviceSynchronize();
    }
    if (e == cudaSuccess && src2) {
        e = cudaMemcpy(stg, src2, bytes, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) {
            launch_pack_q4_0(stg, rows, cols, d_packed, 1, stride, total, 0);
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
    launch_quant_act(d_x, n, d_act, 0);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
extern "C" int b2_op_rmsnorm_quant(const float *d_x, const float *d_w, int64_t n, float eps, int plus_one, void *d_act, float *d_y_opt) {
    ApiScope _scope("b2_op_rmsnorm_quant");
    B2_CHECK(d_x && d_w && d_act && n > 0 && n % QK == 0 && n <= 16384, B2_EINVAL, "b2_op_rmsnorm_quant: n must be a multiple of 32 in (0,16384]");
    launch_rmsnorm_quant(d_x, d_w, n, eps, plus_one, d_act, d_y_opt, 0);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
extern "C" int b2_op_gemv(int type, const void *d_packed, int64_t rows, int64_t cols, const void *d_act, const float *d_bias_or_res,
                          int mode, int act, float *d_y) {
    ApiScope _scope("b2_op_gemv");
    B2_CHECK(type == B2_T_Q4_0, B2_EINVAL, "b2_op_gemv: type %d not supported", type);
    B2_CHECK(d_packed && d_act && d_y && rows >= 0 && cols > 0 && cols % QK == 0, B2_EINVAL, "b2_op_gemv: bad arguments");
    B2_CHECK(mode >= 0 && mode <= 2, B2_EINVAL, "b2_op_gemv: mode %d", mode);
    B2_CHECK(mode != 1 || d_bias_or_res, B2_EINVAL, "b2_op_gemv: mode 1 needs the residual vector");
    B2_CHECK(mode != 2 || rows % 2 == 0, B2_EINVAL, "b2_op_gemv: mode 2 needs an even row count");
    if (rows == 0) return B2_OK;
    launch_gemv_q4_0(d_packed, rows, cols, d_act, d_bias_or_res, mode, act, d_y, 0);
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
extern "C" int b2_op_argmax(const float *d_logits, int64_t n, int32_t *d_out_index) {
    ApiScope _scope("b2_op_argmax");
    B2_CHECK(d_logits && d_out_index && n > 0 && n < (1ll << 31), B2_EINVAL, "b2_op_argmax: bad arguments");
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
