// sampler.cuh — launcher of the top-k sampler (sampler.cu)
#pragma once
#include "common.cuh"
#define SAMPLE_BLOCKS 128
#define SAMPLE_MAX_K 64
uint64_t sample_scratch_bytes(int k);   // zero once before first use; the kernel leaves its counter at zero
// d_ctx / d_ctx_len (or ctx_len_imm) / penalty / last_n: repetition penalty over the last tokens of the context (nullptr / 1.0: off)
int launch_sample_topk(const float *d_logits, int64_t n, int k, float temperature, float top_p, const float *d_u, float u_imm,
                       int32_t *d_out, void *d_scratch, cudaStream_t st, const int *d_ctx = nullptr, const int *d_ctx_len = nullptr,
                       int ctx_len_imm = 0, float penalty = 1.0f, int last_n = 0);
