// prefill.cuh — batched prompt evaluation (the "prompt eval" half of Ollama's /api/generate):
// T prompt tokens go through a weight matrix in ONE pass over its bytes, on the 5th-generation tensor cores
// (tcgen05.mma kind::f16 on bf16 operands, fp32 accumulators in tensor memory), instead of T passes of the decode GEMV.
//
// Numerics are those of the decode path: the activation vector of every token is put on the block
// fixed-point grid (21-bit integers X times 2^(e-19) per 32-element block), the weights are the exact dequantised values.
// Both are split into bf16 pieces WITHOUT rounding: X * 2^(e-19) = sum over three balanced base-128 digits of
// digit * 128^plane * 2^(e-19) (7 significant bits each), a Q4_0 weight d*(q-8) has at most 15 significant bits = two bf16 parts,
// a Q4_K weight (24 bits) three.  Every bf16 x bf16 product is exact in fp32, so the tensor cores add up, in fp32, exactly the
// terms of the dot product of the dequantised row with the gridded activations; only the order and the tensor core's fp32
// addition differ from the CPU oracle.
//
// "PB" = the B operand of the GEMM for a batch of T tokens (T = 16, 32, 64 or 128):
//   [kb 0..nbp)  per 32-element block two images (k-steps of 16 elements) of 3T rows x 16 bf16 in the shared-memory layout
//                tcgen05.mma wants (K-major, no swizzle: 8-row x 16-byte core matrices; row n = plane*T + m for digit plane `plane`
//                of token m): the bf16 of (n, k) at kb*192T + (k>>4)*96T + (n>>3)*256 + ((k&15)>>3)*128 + (n&7)*16 + (k&7)*2.
//                One stage = one cp.async.bulk of two blocks (T <= 64) or one block (T = 128).
#pragma once
#include "kernels.cuh"
#include "layout.cuh"

#define PF_ROWS 128          // weight rows per CTA tile (UMMA M)
#ifndef PF_SEG
#define PF_SEG 8             // stages whose products accumulate in tensor memory before they are drained into fp32 registers (512 k)
#endif
#define PF_THREADS 448       // warps 0-3 accumulator drain + epilogue (TMEM quadrant = warp), 4-11 weight dequantisation, 12 MMA issue, 13 B-operand copies

__host__ __device__ inline size_t pb_block_bytes(int T) { return (size_t)2 * 3 * T * 32; }
__host__ __device__ inline size_t pb_bytes(int T, int64_t n) { return (size_t)nblocks_padded(n) * pb_block_bytes(T); }
__host__ __device__ inline size_t pb_off(int T, int64_t kb, int d, int m, int k) {
    const int n = d * T + m, kk = k & 15;
    return (size_t)kb * pb_block_bytes(T) + (size_t)(k >> 4) * (3 * T * 32) + (size_t)(n >> 3) * 256 + (size_t)(kk >> 3) * 128 + (size_t)(n & 7) * 16 +
           (size_t)(kk & 7) * 2;
}
// smallest supported batch size holding m tokens (0 if m > 128)
inline int pf_batch(int m) { return m <= 16 ? 16 : m <= 32 ? 32 : m <= 64 ? 64 : m <= 128 ? 128 : 0; }
#define PF_MAX_TOKENS 128   // tokens per pass over the weights

enum { PF_IN_NORM = 0, PF_IN_F32 = 2 };
enum { PF_OUT_BIAS = 0, PF_OUT_RES = 1, PF_OUT_ACTMUL = 2 };

// x[M][n] fp32 (row stride ldx) -> PB for T tokens.  in_mode PF_IN_NORM: RMSNorm with norm_w first.
cudaError_t launch_prefill_quant(int in_mode, const float *x, int64_t ldx, int M, int T, int64_t n, const float *norm_w, float eps, int plus_one,
                                 void *pb, cudaStream_t st);
// `part` (may be NULL): scratch of part_floats floats for split-K partial sums (up to 8 * T * rows are used)
// y[m][row] (row stride ldy) = sum_k W[row][k] * grid(x[m])[k]  (+ bias[row] | + res[m][row] | act(gate)*up with ldy = rows/2)
// qt: 0 = Q4_0 matrix, 1 = Q4_K, 2 = Q6_K
// partials_only: the raw split-K sums stay in `part` ([ksplit][T][rows], *ksplit_out pieces, possibly 1) and no epilogue runs: a
// tensor-parallel rank's row-parallel GEMM, whose sums are exchanged before the residual is added
cudaError_t launch_prefill_gemm_qt(int qt, int out_mode, const void *d_packed, int64_t rows, int64_t cols, const void *pb, int M, int T, float *y, int64_t ldy,
                                   const float *bias_or_res, int act, int sm_count, float *part, uint64_t part_floats, cudaStream_t st,
                                   bool partials_only = false, int *ksplit_out = nullptr);
cudaError_t launch_prefill_gemm(int out_mode, const void *d_packed, int64_t rows, int64_t cols, const void *pb, int M, int T, float *y, int64_t ldy,
                                const float *bias_or_res, int act, int sm_count, float *part, uint64_t part_floats, cudaStream_t st);
int pf_pick_ksplit(int n_groups, int sm_count, int kmax);
void launch_prefill_embed(int qt, const void *d_packed, int64_t cols, const int *d_tokens, int M, float scale, float *d_x, cudaStream_t st);
// RoPE (q in place) + KV append of M tokens at pos0.., then causal attention over the cache; out[m][nh*hd] fp32
cudaError_t launch_prefill_attention(float *d_qkv, int64_t ld, int M, int nh, int nkv, int hd, int pos0, const float *d_cos, const float *d_sin,
                                     __half *d_k, __half *d_v, const int32_t *d_page_table, float *d_out, int64_t ldo, cudaStream_t st);
void launch_fill_partials(float *pmax, int *pidx, int n, cudaStream_t st);
