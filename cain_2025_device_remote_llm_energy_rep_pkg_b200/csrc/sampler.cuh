// sampler.cuh — launcher of the top-k sampler (sampler.cu)
#pragma once
#include "common.cuh"
#define SAMPLE_BLOCKS 128
#define SAMPLE_MAX_K 64
uint64_t sample_scratch_bytes(int k);   // zero once before first use; the kernel leaves its counter at zero
int launch_sample_topk(const float *d_logits, int64_t n, int k, float temperature, float top_p, const float *d_u, float u_imm,
                       int32_t *d_out, void *d_scratch, cudaStream_t st);
