/* b2llm.h — C ABI of the B200 generation engine.
 *
 * This is the drop-in boundary for the one path the reference spends its time on: everything that
 * happens inside Ollama's `POST /api/generate` (reference call site experiment/RunnerConfig.py:128-131,
 * install note README.md:29-31; stage list SURVEY.md §3.3).  The reference binds that path over HTTP,
 * not over an FFI, so there is no reference header to mirror; the entry points below are the calls the
 * Ollama-compatible server (server.py) and the runner plugin (runner/nvml_energy.py) make through
 * ctypes, one per stage of that request, plus op-level entry points the parity tests drive.
 *
 * Conventions: every function returns 0 on success and a negative B2_E* code on failure;
 * b2_last_error() returns a thread-local message for the last failure on the calling thread.
 * Handles are opaque.  The engine owns all device memory; the caller owns all host buffers.
 * Pointers named `d_*` are DEVICE pointers (the op-level entry points), everything else is host.
 * One session = one request; a session is not re-entrant; sessions on different engines are
 * independent.  Nothing here falls back to the CPU: without a CUDA device every compute call fails.
 */
#ifndef B2LLM_H
#define B2LLM_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_OK            0
#define B2_EINVAL       -1   /* bad argument / unsupported shape */
#define B2_ECUDA        -2   /* CUDA runtime error (message has the CUDA string) */
#define B2_ENOMEM       -3
#define B2_ECTX         -4   /* context window full */
#define B2_ENVML        -5   /* NVML unavailable or failed */
#define B2_ENCCL        -6

/* GGML tensor types understood by the engine (same numeric values as ggml/gguf).  Matrices: Q4_0 and Q4_K for the layers
 * (the parts of a fused QKV / gate-up matrix share one type), Q4_0 / Q4_K / Q6_K for token_embd / output; vectors: F32. */
#define B2_T_F32   0
#define B2_T_Q4_0  2
#define B2_T_Q8_0  8
#define B2_T_Q4_K  12
#define B2_T_Q6_K  14

#define B2_ACT_SILU       0
#define B2_ACT_GELU_TANH  1

typedef struct b2_engine  b2_engine;
typedef struct b2_model   b2_model;
typedef struct b2_session b2_session;

/* Architecture record: the hyper-parameters of one of the seven model tags the reference sweeps
 * (experiment/RunnerConfig.py:80).  Mirrors oracle/archs.py:Arch. */
typedef struct b2_arch_cfg {
    int32_t hidden;          /* H */
    int32_t n_layers;        /* L */
    int32_t n_heads;         /* query heads */
    int32_t n_kv_heads;
    int32_t head_dim;        /* 32-multiple, <= 256 */
    int32_t ffn;             /* I */
    int32_t vocab;           /* V */
    int32_t tied;            /* LM head = token embedding */
    int32_t qkv_bias;        /* qwen2 */
    int32_t act;             /* B2_ACT_* */
    int32_t norm_plus_one;   /* gemma: x_hat * (1 + w) */
    int32_t embed_scale;     /* gemma: embedding * sqrt(H) */
    int32_t rope_llama3;     /* 1: apply llama-3.1 frequency scaling with the 4 values below */
    int32_t max_ctx;         /* context window the KV cache is sized for */
    float   rope_theta;
    float   rms_eps;
    float   rope_factor, rope_low_freq, rope_high_freq, rope_orig_ctx;
} b2_arch_cfg;

/* One tensor of a model, GGUF naming ("token_embd.weight", "blk.3.attn_q.weight", ...), host bytes
 * in GGML block layout (row-major rows x cols, row = cols/block * block_bytes). */
typedef struct b2_tensor_desc {
    const char *name;
    int32_t     type;        /* B2_T_* */
    int64_t     rows, cols;  /* cols == 0: a vector of `rows` f32 */
    const void *data;
    uint64_t    nbytes;
} b2_tensor_desc;

typedef struct b2_sampler_cfg {
    int32_t  top_k;          /* 0 or 1: greedy argmax (lowest index wins ties); 2..64: sample among the k largest logits */
    float    temperature;    /* used when top_k > 1; must be > 0 */
    float    top_p;          /* (0,1]; 1.0 disables the nucleus cut */
    uint64_t seed;           /* counter-based: draw i of a session (since create/reset) uses the uniform of (seed, i) */
    float    repeat_penalty; /* used when top_k > 1.  0 or 1: off.  Ollama's options.repeat_penalty (llama.cpp's rule): the logit of every distinct
                              * token among the last repeat_last_n tokens of the context (prompt included) is divided by it when positive,
                              * multiplied by it otherwise, before the candidates are selected */
    int32_t  repeat_last_n;  /* window of the penalty in tokens (Ollama's default 64); 0: off */
} b2_sampler_cfg;

typedef struct b2_stats {
    uint64_t prompt_eval_count, eval_count;      /* tokens */
    uint64_t prompt_eval_ns, eval_ns;            /* device time, CUDA events */
    uint64_t kernel_launches;                    /* kernels launched by this session so far */
} b2_stats;

const char *b2_last_error(void);
const char *b2_version(void);

/* ---- engine / model / session: the stages of one /api/generate request -------------------------- */
/* Replaces: Ollama runner start-up (README.md:29-31 "install Ollama").  tp_size 1 = single GPU.
 * tp_size 2, 4 or 8: this engine is rank tp_rank of a tensor-parallel group, one process and one GPU per rank (SURVEY.md §8d
 * config 4 / §8e; no reference counterpart: Ollama places a model on one GPU here).  The ranks exchange partial sums through
 * memory they map into each other (see b2_engine_tp_connect), not through a communicator. */
int b2_engine_create(int device_ordinal, int tp_rank, int tp_size, b2_engine **out);
int b2_engine_destroy(b2_engine *e);
int b2_engine_device(const b2_engine *e, int *device_ordinal);
/* Tensor-parallel set-up, before any model is created on the engine: every rank calls b2_engine_tp_handle (B2_TP_HANDLE_BYTES
 * bytes out: a CUDA IPC handle of its exchange block), the host side gathers the handles of all ranks in rank order (any
 * transport: torch.distributed, a pipe, a file), and every rank passes the tp_size * B2_TP_HANDLE_BYTES bytes to
 * b2_engine_tp_connect, which maps the peers' blocks (needs peer access between the GPUs: NVLink / NVSwitch on a B200 box).
 * From then on every rank must make the same model / session / prefill / decode calls with the same arguments; each rank holds
 * 1/tp_size of every layer matrix and returns the same tokens and logits (the LM head is multiplied row-parallel in the steps of
 * a greedy b2_decode call, with the arg-max exchanged between the ranks, and in full wherever logits are handed out).  A rank whose peer does not show up
 * within 2 s fails the call with B2_ENCCL instead of hanging.  Prompts of three tokens and more are evaluated in
 * batches in this mode too (the partial sums of the row-parallel prompt GEMMs travel the same way, a batch at a time). */
#define B2_TP_HANDLE_BYTES 64
int b2_engine_tp_handle(b2_engine *e, void *handle_out);
int b2_engine_tp_connect(b2_engine *e, const void *handles);

/* Replaces: Ollama model load on first request for a tag (RunnerConfig.py:128, field "model").
 * Uploads the tensors once and re-packs them into the engine's HBM layout. */
int b2_model_create(b2_engine *e, const b2_arch_cfg *cfg, const b2_tensor_desc *tensors, int n_tensors, b2_model **out);
/* Same model, but the deterministic random-init weights (oracle/randinit.py recipe) are generated and
 * quantised on the device; byte-identical to b2_randinit_tensor + b2_model_create for Q4_0.  qtype/head_qtype: B2_T_*;
 * Q4_K / Q6_K get synthetic blocks with the same weight statistics (benchmarks only, no host counterpart). */
int b2_model_create_random(b2_engine *e, const b2_arch_cfg *cfg, uint64_t seed, int qtype, int head_qtype, b2_model **out);
int b2_model_destroy(b2_model *m);
/* Bytes of matmul weights one decoded token reads (the roofline's algorithmic bytes). */
int b2_model_weight_bytes(const b2_model *m, uint64_t *out);

int b2_session_create(b2_model *m, int max_ctx, b2_session **out);
int b2_session_reset(b2_session *s);
int b2_session_destroy(b2_session *s);
int b2_session_pos(const b2_session *s, int *pos);
int b2_session_stats(const b2_session *s, b2_stats *out);

/* Replaces: prompt evaluation inside /api/generate ("prompt_eval_count"/"prompt_eval_duration").
 * tokens: host int32[n]. logits_out: host float[vocab] for the last prompt token, or NULL. */
int b2_prefill(b2_session *s, const int32_t *tokens, int n, float *logits_out);
/* Replaces: the decode loop inside /api/generate ("eval_count"/"eval_duration").  Runs n_steps
 * sample->forward steps starting from the logits left by prefill/decode.  out_tokens: host
 * int32[n_steps].  logits_out: host float[n_steps*vocab] (the logits each token was sampled from) or NULL. */
int b2_decode(b2_session *s, int n_steps, const b2_sampler_cfg *sampler, int32_t *out_tokens, float *logits_out);
/* Teacher-forced variant for parity tests: feeds the given tokens, returns logits after each. */
int b2_forward_tokens(b2_session *s, const int32_t *tokens, int n, float *logits_out /* n*vocab or NULL */);

/* ---- measurement ------------------------------------------------------------------------------- */
/* New (no reference counterpart; BASELINE.json north_star): nvmlDeviceGetTotalEnergyConsumption in mJ. */
int b2_nvml_energy_mj(int device_ordinal, uint64_t *out_mj);
int b2_nvml_power_mw(int device_ordinal, uint32_t *out_mw);
int b2_nvml_util(int device_ordinal, uint32_t *gpu_pct, uint32_t *mem_pct);
int b2_nvml_clocks(int device_ordinal, uint32_t *sm_mhz, uint32_t *sm_max_mhz, uint64_t *throttle_reasons);

/* ---- host-side deterministic random init (input generator, oracle/randinit.py recipe) ----------- */
/* Fills `out` with rows [row0,row0+nrows) of the tensor: GGML blocks of `type`, or f32 when cols==0/type F32. */
int b2_randinit_tensor(uint64_t seed, uint32_t tensor_id, int64_t rows, int64_t cols, int type,
                       float sigma, float offset, int64_t row0, int64_t nrows, void *out, uint64_t out_bytes);

/* ---- device memory helpers for callers without a CUDA binding (tests drive ops through these) ---- */
int b2_dev_alloc(uint64_t nbytes, void **d_out);
int b2_dev_free(void *d_ptr);
int b2_dev_upload(void *d_dst, const void *src, uint64_t nbytes);
int b2_dev_download(void *dst, const void *d_src, uint64_t nbytes);
int b2_dev_sync(void);

/* ---- op-level entry points (device pointers; default stream; used by tests/bench/profiling) ------ */
/* Weight re-pack: GGML row-major blocks -> engine layout.  d_packed must hold
 * b2_packed_bytes(type, rows, cols) bytes. interleave2: rows 2i/2i+1 of the packed matrix are row i of
 * `src` and row i of `src2` (gate/up pairing); src2 NULL otherwise. */
uint64_t b2_packed_bytes(int type, int64_t rows, int64_t cols);
int b2_op_pack_weights(int type, int64_t rows, int64_t cols, const void *src, const void *src2, void *d_packed);
/* Activation grid: x f32[n] -> three int8 digit planes + per-32 scale (b2_act_bytes(n) bytes; layout in DESIGN.md). */
uint64_t b2_act_bytes(int64_t n);
int b2_op_quant_act(const float *d_x, int64_t n, void *d_act);
int b2_op_rmsnorm_quant(const float *d_x, const float *d_w, int64_t n, float eps, int plus_one, void *d_act, float *d_y_opt);
/* y[rows] = W[rows,cols] * act (+ bias) ; mode 0: y = Wx(+bias); 1: y = res + Wx; 2: rows are (gate,up)
 * pairs, y[rows/2] = act(gate)*up. */
int b2_op_gemv(int type, const void *d_packed, int64_t rows, int64_t cols, const void *d_act,
               const float *d_bias_or_res, int mode, int act, float *d_y);
/* Same product with an f32 input vector: the kernel puts x on the activation grid itself (the engine's down projection
 * consumes act(gate)*up this way).  mode 0 (+bias) or 1 (+residual). */
int b2_op_gemv_f32(int type, const void *d_packed, int64_t rows, int64_t cols, const float *d_x, const float *d_bias_or_res,
                   int mode, float *d_y);
/* The engine's fused form: RMSNorm(x, norm_w) -> activation grid -> y = W * (.) with the same epilogues; mode 3 writes
 * logits to d_y and the arg-max index (lowest index on ties) to d_argmax; mode 2 can also return the activation-grid
 * image of its result in d_act_out (b2_act_bytes(rows/2) bytes, may be NULL). */
int b2_op_norm_gemv(int type, const void *d_packed, int64_t rows, int64_t cols, const float *d_x, const float *d_norm_w, float eps,
                    int plus_one, const float *d_bias_or_res, int mode, int act, float *d_y, void *d_act_out, int32_t *d_argmax);
/* Batched form of b2_op_norm_gemv / b2_op_gemv_f32 for M prompt tokens: one pass over the weights on the tcgen05 tensor cores
 * (prompt evaluation of Ollama's /api/generate; no reference file).  d_x [M][cols], d_norm_w NULL = no
 * RMSNorm; mode 0: + d_bias_or_res[rows] (may be NULL); 1: + d_bias_or_res[M][rows]; 2: act(gate)*up, d_y is [M][rows/2]. */
int b2_op_prefill_gemm(int type, const void *d_packed, int64_t rows, int64_t cols, const float *d_x, int M, const float *d_norm_w,
                       float eps, int plus_one, const float *d_bias_or_res, int mode, int act, float *d_y);
/* RoPE (rotate-half) on q[nh,hd] in place and on k, then f16 append of k,v at position pos of a paged cache. */
int b2_op_rope_kv(float *d_qkv, int nh, int nkv, int hd, int pos, const float *d_cos, const float *d_sin,
                  void *d_kcache, void *d_vcache, const int32_t *d_page_table, int page_tokens);
/* softmax(q k^T / sqrt(hd)) v over positions [0,n_ctx) of the paged f16 cache; out f32[nh*hd]. */
int b2_op_attn_decode(const float *d_q, int nh, int nkv, int hd, int n_ctx, const void *d_kcache, const void *d_vcache,
                      const int32_t *d_page_table, int page_tokens, float *d_out);
/* The engine's fused form of the two calls above: rotates q and the new k from the UN-rotated QKV output, appends k,v at
 * `pos`, attends over positions [0,pos], and writes the result as f32 (d_out, may be NULL) and/or on the activation grid
 * (d_act_out, b2_act_bytes(nh*hd) bytes, may be NULL).  max_ctx sizes the kernel's shared-memory score buffer. */
int b2_op_attn_fused(const float *d_qkv, int nh, int nkv, int hd, int pos, const float *d_cos, const float *d_sin, void *d_kcache,
                     void *d_vcache, const int32_t *d_page_table, int page_tokens, int max_ctx, float *d_out, void *d_act_out);
/* Greedy sampler: index of the maximum (lowest index on ties) of f32[n]. */
int b2_op_argmax(const float *d_logits, int64_t n, int32_t *d_out_index);
/* No counterpart in the reference.  Same result as b2_op_argmax, for callers that must not allocate device memory: the
 * sampling thread of a server while a measured request is decoding (cudaMalloc/cudaFree stall the whole device and would
 * end up in the recorded time and energy) and processes that have already filled the card with weights and KV cache.
 * The scratch is a 2 KB array inside the library image, present on every GPU that loads the library until exit; the
 * call uses no cudaMalloc, cudaFree or allocating library routine.  n may be at most B2_ARGMAX_NOALLOC_MAX_N; larger
 * inputs are refused with B2_EINVAL.  The first call on a device looks the array up (cudaGetSymbolAddress, with that
 * device current): make it before any stream capture begins.  All calls on a device share the array: calls whose GPU
 * work overlaps (different streams or threads) race; this entry point runs on the default stream, asynchronously. */
#define B2_ARGMAX_NOALLOC_MAX_N 1048576
int b2_op_argmax_noalloc(const float *d_logits, int64_t n, int32_t *d_out_index);
/* Top-k / temperature / top-p sampler with a caller-supplied uniform u in [0,1). */
int b2_op_sample_topk(const float *d_logits, int64_t n, int top_k, float temperature, float top_p, float u, int32_t *d_out_index);
/* The same with the repetition penalty of b2_sampler_cfg over the ctx_len tokens in d_ctx (device int32). */
int b2_op_sample_topk_penalty(const float *d_logits, int64_t n, int top_k, float temperature, float top_p, float u, const int32_t *d_ctx,
                              int ctx_len, float repeat_penalty, int repeat_last_n, int32_t *d_out_index);

#ifdef __cplusplus
}
#endif
#endif /* B2LLM_H */
