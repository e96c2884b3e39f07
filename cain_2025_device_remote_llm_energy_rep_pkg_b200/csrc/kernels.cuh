// kernels.cuh — launchers of the decode-path kernels (all asynchronous on `st`).
// Each mirrors one stage of Ollama's /api/generate decode loop.
#pragma once
#include "common.cuh"

#ifndef ATTN_NSPLIT
#define ATTN_NSPLIT 8       // KV splits per head in the decode attention kernel
#endif
#ifndef ATTN_CTA_CTX
#define ATTN_CTA_CTX 256    // up to this many positions one 1024-thread CTA per head (eight groups merging through shared memory, two pages of
                            // keys per group in flight): +1.3 % tokens/s over the cluster kernel at contexts 32..288 (it occupies nh SMs instead of
                            // all of them, so the O projection's CTAs become resident while it runs), -3.3 % at contexts up to 1 365
#endif
#define ATTN_VARIANTS 3
#ifndef ATTN_SHORT_CTX
#define ATTN_SHORT_CTX 32   // up to this many positions the fused attention runs one CTA per head (no cluster); measured on llama3.1:8b,
                            // contexts 32..288: 96 -> 659, 64 / 32 / 16 -> 665 tokens/s, 160 -> 641 (the cluster variant wins from the first page pair on)
#endif
#define ARGMAX_BLOCKS 128   // stage-1 blocks of the greedy sampler

// device-resident request state: lets a captured CUDA graph be replayed for every token
struct DevState {
    int pos;        // tokens already in the KV cache
    int step;       // decode step counter (index into out_tokens)
    int cur_token;  // token fed to the next forward
    int prompt_idx; // next prompt token to feed during prefill
};

// weight re-pack / init into the streaming layout (layout.cuh)
// src_cols / src_col0 (0 / 0 = the whole row): the source rows are src_cols wide and columns [src_col0, src_col0 + cols) are taken;
// src_row0 of the initialisers: first row of the window inside the whole tensor (tensor-parallel ranks fill their share).
void launch_pack_q4_0(const void *d_src, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                      int64_t src_cols = 0, int64_t src_col0 = 0);
void launch_pack_q4_k(const void *d_src, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                      int64_t src_cols = 0, int64_t src_col0 = 0);
void launch_pack_q6_k(const void *d_src, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                      int64_t src_cols = 0, int64_t src_col0 = 0);
void launch_randinit_q4_0(uint64_t key, float sigma, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                          int64_t src_row0 = 0, int64_t src_cols = 0, int64_t src_col0 = 0);
void launch_randinit_q4_k(uint64_t key, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                          int64_t src_row0 = 0, int64_t src_cols = 0, int64_t src_col0 = 0);
void launch_randinit_q6_k(uint64_t key, int64_t rows, int64_t cols, void *d_packed, int64_t dst_row0, int64_t dst_row_stride, cudaStream_t st,
                          int64_t src_row0 = 0, int64_t src_cols = 0, int64_t src_col0 = 0);
void launch_randinit_f32(uint64_t key, float sigma, float offset, int64_t n, float *d_out, cudaStream_t st, int64_t i0 = 0);

// matrix format index of the kernels: 0 = Q4_0, 1 = Q4_K, 2 = Q6_K
static inline int qt_of(int type) { return type == B2_T_Q6_K ? 2 : type == B2_T_Q4_K ? 1 : 0; }
// streaming GEMV (gemv.cu)
enum { GEMV_IN_NORM = 0, GEMV_IN_ACT = 1, GEMV_IN_F32 = 2 };
enum { GEMV_OUT_BIAS = 0, GEMV_OUT_RES = 1, GEMV_OUT_ACTMUL = 2, GEMV_OUT_LOGITS = 3 };
int gemv_grid(int n_tiles, int tile_align);
cudaError_t launch_gemv(int in_mode, int out_mode, const void *d_packed, int64_t rows, int64_t cols, const float *x, const float *norm_w,
                        float eps, int plus_one, const void *act_in, float *y, const float *bias_or_res, void *act_out, int act,
                        float *pmax, int *pidx, int *grid_out, const void *next_w, uint64_t next_bytes, const float *xsq_in, int xsq_n,
                        float *xsq_out, cudaStream_t st, bool pdl);
cudaError_t launch_gemv_qt(int qt, int in_mode, int out_mode, const void *d_packed, int64_t rows, int64_t cols, const float *x, const float *norm_w,
                           float eps, int plus_one, const void *act_in, float *y, const float *bias_or_res, void *act_out, int act,
                           float *pmax, int *pidx, int *grid_out, const void *next_w, uint64_t next_bytes, const float *xsq_in, int xsq_n,
                           float *xsq_out, cudaStream_t st, bool pdl);
// tensor-parallel modes (gemv_tp.cu): GEMV_IN_TPNORM consumes an exchange round, GEMV_OUT_TPPART produces one
enum { GEMV_IN_TPNORM = 3, GEMV_OUT_TPPART = 4 };
struct GemvTp {
    int n = 1;
    void *dst[8] = {};                     // OUT_TPPART: slot of this rank in every rank's exchange block of the round ((value, tag) pairs)
    const float *res = nullptr;            // IN_TPNORM: residual before the exchange,
    const void *slots = nullptr;           //            local slots of the round (slot r at r * slot_stride pairs),
    int slot_stride = 0;
    const unsigned long long *fwd = nullptr;   // forward passes completed on this engine (device counter)
    int round = 0, rounds = 1;                 // tag of the round: low 32 bits of *fwd * rounds + round + 1
    float *xout = nullptr;                 //            new residual (written by CTA 0)
    unsigned int *err = nullptr;
};
cudaError_t launch_gemv_tp(const GemvTp &tp, int qt, int in_mode, int out_mode, const void *d_packed, int64_t rows, int64_t cols,
                           const float *norm_w, float eps, int plus_one, const void *act_in, float *y, const float *bias, int act,
                           float *pmax, int *pidx, int *grid_out, cudaStream_t st, bool pdl, int row_off = 0);
// greedy token of a tensor-parallel group: local arg-max partials -> one tagged 16-byte word per rank -> the same index everywhere
void launch_tp_argmax(const float *pmax, const int *pidx, int n, void *const *peer_slots, int tp_n, int rank, const void *local_slots,
                      const unsigned long long *fwd, unsigned int *err, int32_t *d_out, cudaStream_t st);
// batched prompt evaluation on a tensor-parallel rank: split-K partials of a row-parallel GEMM -> tagged pairs into every rank's batch
// buffer (peer_slots[p] = slot of this rank on rank p), then y = res + sum over the ranks' slots of this GPU's buffer (rank order)
void launch_pf_tp_exchange(const float *part, int ksplit, int T, int M, int rows, void *const *peer_slots, const void *local_slots, size_t slot_stride,
                           int tp_n, const float *res, float *y, int64_t ld, const unsigned long long *fwd, int rounds, int round, unsigned int *err,
                           cudaStream_t st);
void launch_tp_close(const void *slots, int slot_stride, int n, int n_vec, unsigned long long *fwd, int rounds, int round, bool wait, unsigned int *err,
                     cudaStream_t st);
void launch_embed_row(int qt, const void *d_packed, int64_t cols, const int *d_token, float scale, float *d_x, float *d_xsq, cudaStream_t st);
void launch_sumsq(const float *d_x, int64_t n, float *d_xsq, cudaStream_t st);

// small stages (kernels.cu)
void launch_quant_act(const float *d_x, int64_t n, void *d_act, cudaStream_t st);
void launch_rmsnorm_quant(const float *d_x, const float *d_w, int64_t n, float eps, int plus_one, void *d_act, float *d_y_opt, cudaStream_t st);
void launch_rope_kv(float *d_qkv, int nh, int nkv, int hd, const int *d_pos, int pos_imm, const float *d_cos, const float *d_sin,
                    __half *d_k, __half *d_v, const int32_t *d_page_table, cudaStream_t st);
// scratch: float[nh*ATTN_NSPLIT*(hd+2)] then int counters[nh] (zeroed once; the kernel leaves them zero)
uint64_t attn_scratch_bytes(int nh, int hd);
void launch_attn_decode(const float *d_q, int nh, int nkv, int hd, const int *d_pos, int nctx_imm, const __half *d_k, const __half *d_v,
                        const int32_t *d_page_table, float *d_out, void *d_scratch, int max_ctx, cudaStream_t st);
// scratch: float[ARGMAX_BLOCKS] + int[ARGMAX_BLOCKS] + int counter (zeroed once)
uint64_t argmax_scratch_bytes();
void launch_argmax(const float *d_logits, int64_t n, int32_t *d_out, void *d_scratch, cudaStream_t st);
// final arg-max over the per-CTA partials of the LM-head GEMV, token commit and next-token bookkeeping
void launch_argmax_partials(const float *pmax, const int *pidx, int n, int32_t *d_out, cudaStream_t st);
// fused RoPE + KV append + paged attention + activation grid (attn.cu); one cluster of ATTN_NSPLIT CTAs per head
cudaError_t launch_attn_fused(const float *d_qkv, int nh, int nkv, int hd, const int *d_pos, int pos_imm, const float *d_cos,
                              const float *d_sin, __half *d_k, __half *d_v, const int32_t *d_page_table, float *d_out, void *d_act_out,
                              int max_ctx, cudaStream_t st, bool pdl);
