// prefill.cuh — batched prompt evaluation (the "prompt eval" half of Ollama's /api/generate):
// T prompt tokens go through a weight matrix in ONE pass over its bytes, on the 5th-generation tensor cores
// (tcgen05.mma kind::i8, accumulators in tensor memory), instead of T passes of the decode GEMV.
//
// Numerics are those of the decode path: the activation vector of every token is put on the block
// fixed-point grid (21-bit integers as three balanced base-128 int8 digits), the weights are the int4 values times 16, so
// each 32-element block contributes three exact int32 dot products per (row, token); they are combined and scaled by
// d_w(row, block) * s_act(token, block) in fp32.
//
// "PB" = the B operand of the GEMM for a batch of T tokens (T = 16, 32 or 64; N = 3T int8 columns):
//   [kb 0..nbp)  one image of N rows x 32 bytes per 32-element block, already in the shared-memory layout tcgen05.mma wants
//                (K-major, no swizzle: 8-row x 16-byte core matrices; row n = d*T + m for digit plane d of token m):
//                byte of (n, k) at kb*N*32 + (n>>3)*256 + (k>>4)*128 + (n&7)*16 + (k&15)  -- one cp.async.bulk per stage
//   then per block kb 2T floats: sa[T] = 2^(e-19)/16 (the 1/16 undoes the weights' x16; 0 for padding blocks and tokens >= M)
//   and ts[T] = the sum of the token's gridded activations over the block (the Q4_K "- dmin*m" term multiplies it).
#pragma once
#include "kernels.cuh"
#include "layout.cuh"

#define PF_ROWS 128          // weight rows per CTA tile (UMMA M)
#define PF_STAGES 3          // shared-memory stages of one block-group (4 blocks = 128 k) each
#define PF_THREADS 448       // warps 0-7 epilogue (TMEM quadrant = warp % 4, token half = warp / 4), 8-11 weight unpack, 12 MMA issue, 13 B-operand copies

__host__ __device__ inline size_t pb_block_bytes(int T) { return (size_t)3 * T * 32; }
__host__ __device__ inline size_t pb_scale_off(int T, int64_t nbp) { return (size_t)nbp * pb_block_bytes(T); }
__host__ __device__ inline size_t pb_bytes(int T, int64_t n) {
    const int64_t nbp = nblocks_padded(n);
    return pb_scale_off(T, nbp) + (size_t)nbp * T * 8;
}
__host__ __device__ inline size_t pb_off(int T, int64_t kb, int d, int m, int k) {
    const int n = d * T + m;
    return (size_t)kb * pb_block_bytes(T) + (size_t)(n >> 3) * 256 + (size_t)(k >> 4) * 128 + (size_t)(n & 7) * 16 + (size_t)(k & 15);
}
// smallest supported batch size holding m tokens (0 if m > 64)
inline int pf_batch(int m) { return m <= 16 ? 16 : m <= 32 ? 32 : m <= 64 ? 64 : 0; }

enum { PF_IN_NORM = 0, PF_IN_F32 = 2 };
enum { PF_OUT_BIAS = 0, PF_OUT_RES = 1, PF_OUT_ACTMUL = 2 };

// x[M][n] fp32 (row stride ldx) -> PB for T tokens.  in_mode PF_IN_NORM: RMSNorm with norm_w first.
cudaError_t launch_prefill_quant(int in_mode, const float *x, int64_t ldx, int M, int T, int64_t n, const float *norm_w, float eps, int plus_one,
                                 void *pb, cudaStream_t st);
// `part` (may be NULL): scratch of part_floats floats for split-K partial sums (up to 8 * T * rows are used)
// y[m][row] (row stride ldy) = sum_k W[row][k] * grid(x[m])[k]  (+ bias[row] | + res[m][row] | act(gate)*up with ldy = rows/2)
// qt: 0 = Q4_0 matrix, 1 = Q4_K matrix
cudaError_t launch_prefill_gemm_qt(int qt, int out_mode, const void *d_packed, int64_t rows, int64_t cols, const void *pb, int M, int T, float *y, int64_t ldy,
                                   const float *bias_or_res, int act, int sm_count, float *part, uint64_t part_floats, cudaStream_t st);
cudaError_t launch_prefill_gemm(int out_mode, const void *d_packed, int64_t rows, int64_t cols, const void *pb, int M, int T, float *y, int64_t ldy,
                                const float *bias_or_res, int act, int sm_count, float *part, uint64_t part_floats, cudaStream_t st);
void launch_prefill_embed(int qt, const void *d_packed, int64_t cols, const int *d_tokens, int M, float scale, float *d_x, cudaStream_t st);
// RoPE (q in place) + KV append of M tokens at pos0.., then causal attention over the cache; out[m][nh*hd] fp32
cudaError_t launch_prefill_attention(float *d_qkv, int64_t ld, int M, int nh, int nkv, int hd, int pos0, const float *d_cos, const float *d_sin,
                                     __half *d_k, __half *d_v, const int32_t *d_page_table, float *d_out, int64_t ldo, cudaStream_t st);
void launch_fill_partials(float *pmax, int *pidx, int n, cudaStream_t st);
