// engine.cu — model / session objects and the C ABI of libb2llm (include/b2llm.h).
// Replaces what happens inside Ollama's POST /api/generate (reference call site
// experiment/RunnerConfig.py:128-131): model load, prompt evaluation, the batch-1 decode loop.
#include "kernels.cuh"
#include "prefill.cuh"
#include "sampler.cuh"
#include "layout.cuh"
#include "randinit.cuh"
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <algorithm>
#include <vector>
#include <string>
#include <map>

// ------------------------------------------------------------------------------------ errors
static thread_local char g_err[1024] = "";
void b2_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}
extern "C" const char *b2_last_error(void) { return g_err; }
static bool debug_stale() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("B2_DEBUG_STALE"); v = (e && e[0] == '1') ? 1 : 0; }
    return v == 1;
}
ApiScope::ApiScope(const char *n) : name(n) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess && debug_stale()) fprintf(stderr, "[b2llm] stale CUDA error on entry to %s: %s\n", n, cudaGetErrorString(e));
}
ApiScope::~ApiScope() {
    if (!debug_stale()) return;
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) fprintf(stderr, "[b2llm] CUDA error pending on exit from %s: %s\n", name, cudaGetErrorString(e));
}
extern "C" const char *b2_version(void) { return "b2llm 0.1 (sm_100a)"; }

// ------------------------------------------------------------------------------------ objects
// Exchange memory of a tensor-parallel group (one engine = one rank = one process = one GPU).  Every rank allocates one block,
// hands its CUDA IPC handle to the others (b2_engine_tp_handle / b2_engine_tp_connect) and maps theirs.  Layout of a block:
//   pair  slots[2 buffers][8 source ranks][TP_MAX_VEC]   pair = (f32 partial sum, u32 tag of the exchange round), stored by the
//                                                        peers (and by this rank, into its own slot) from inside their GEMVs
//   u64   forwards                                       forward passes this rank has completed (the tags derive from it)
//   u32   error                                          1 + a rank whose pairs did not arrive in time (sticky)
//   u4    amax[8 source ranks]                           (best logit, its index, tag, tag) of every rank's share of the vocabulary
// Buffer 0 carries the O-projection rounds, buffer 1 the down-projection rounds, so a rank that is one exchange ahead never
// overwrites slots a slower rank still reads (a rank cannot be two exchanges ahead: each exchange needs every rank's pairs).
#define TP_MAX_VEC 8192
#define TP_MAX_RANKS 8
#define TP_PF_TOKENS 64
struct TpLink {
    int n = 1, rank = 0;
    void *local = nullptr;
    void *peer[TP_MAX_RANKS] = {};
    bool connected = false;
    static constexpr size_t slots_bytes() { return (size_t)2 * TP_MAX_RANKS * TP_MAX_VEC * 8; }
    // batched prompt evaluation: the same two buffers of eight slots, TP_PF_TOKENS tokens deep (behind the 512 bytes of flags)
    static constexpr size_t batch_slot_pairs() { return (size_t)TP_PF_TOKENS * TP_MAX_VEC; }
    static constexpr size_t bytes() { return slots_bytes() + 512 + (size_t)2 * TP_MAX_RANKS * batch_slot_pairs() * 8; }
    void *batch_slot(int on_rank, int buf, int src) const {
        return (char *)peer[on_rank] + slots_bytes() + 512 + ((size_t)buf * TP_MAX_RANKS + src) * batch_slot_pairs() * 8;
    }
    void *amax(int on_rank) const { return (char *)peer[on_rank] + slots_bytes() + 256; }
    void *slot(int on_rank, int buf, int src) const { return (char *)peer[on_rank] + ((size_t)buf * TP_MAX_RANKS + src) * TP_MAX_VEC * 8; }
    unsigned long long *forwards() const { return (unsigned long long *)((char *)local + slots_bytes()); }
    unsigned int *error() const { return (unsigned int *)((char *)local + slots_bytes() + 128); }
};
struct b2_engine {
    int device = 0, tp_rank = 0, tp_size = 1;
    cudaStream_t stream = nullptr;
    int sm_count = 0;
    TpLink tp;
};
struct Mat {
    void *packed = nullptr;
    int64_t rows = 0, cols = 0;
    int type = B2_T_Q4_0;
};
struct LayerW {
    float *attn_norm = nullptr, *ffn_norm = nullptr, *qkv_bias = nullptr;
    Mat qkv, o, gateup, down;
    // attn_q / attn_k / attn_v in different formats (llama.cpp's Q4_K_M keeps attn_v in Q6_K): three matrices, three GEMVs
    bool split_qkv = false;
    Mat qkv_part[3];
};
struct LayerTypes { int q, k, v, o, gateup, down; };
struct ModelTypes {
    int embd, head;
    std::vector<LayerTypes> layers;
    static ModelTypes uniform(int n_layers, int qtype, int head_qtype) {
        ModelTypes t;
        t.embd = t.head = head_qtype;
        t.layers.assign(n_layers, LayerTypes{qtype, qtype, qtype, qtype, qtype, qtype});
        return t;
    }
};
struct b2_model {
    b2_engine *e = nullptr;
    b2_arch_cfg cfg{};
    Mat embd, head;
    float *out_norm = nullptr;
    std::vector<LayerW> layers;
    float *rope_cos = nullptr, *rope_sin = nullptr;
    uint64_t weight_bytes = 0;
    std::vector<void *> allocs;
    // a rank of a tensor-parallel group holds n_heads / tp query heads, n_kv_heads / tp KV heads and ffn / tp rows of gate / up
    // (column-parallel QKV, gate, up; row-parallel O, down); norms, embedding and LM head are replicated.  tp = 1: everything.
    int tp = 1, tp_rank = 0;
    int heads() const { return cfg.n_heads / tp; }
    int kv_heads() const { return cfg.n_kv_heads / tp; }
    int ffn() const { return cfg.ffn / tp; }
    int q_dim() const { return heads() * cfg.head_dim; }
    int kv_dim() const { return kv_heads() * cfg.head_dim; }
};
enum { G_PREFILL_BODY = 0, G_PREFILL_LAST, G_DECODE, G_DECODE_LOG, G_FORCED, G_SAMPLE, G_SAMPLE_LOG, G_DECODE_VP, G_COUNT };   // _VP: tensor parallel, LM head split by rows
struct b2_session {
    b2_model *m = nullptr;
    int max_ctx = 0, n_pages = 0;
    __half *kcache = nullptr, *vcache = nullptr;   // [L][n_pages][nkv][PAGE_TOKENS][hd]
    int32_t *page_table = nullptr;
    DevState *state = nullptr;
    int *d_prompt = nullptr;
    int prompt_cap = 0;
    int *d_out_tokens = nullptr;
    int out_cap = 0;
    int *d_sampled = nullptr;
    int *d_ctx = nullptr;          // the token at every position of the context (the repetition penalty looks back over it)
    float *x = nullptr, *qkv = nullptr, *attn_out = nullptr, *h = nullptr, *logits = nullptr;
    void *act_attn = nullptr, *act_h = nullptr, *attn_scratch = nullptr, *argmax_scratch = nullptr;
    float *pmax = nullptr;
    int *pidx = nullptr;
    float *xsq = nullptr;          // per-CTA partial sums of squares of the residual stream (input of the next RMSNorm)
    int head_grid = 0;
    bool use_pdl = true;
    bool fused_attn = true;
    float *logits_log = nullptr;
    uint64_t logits_log_cap = 0;   // in floats
    // top-k / top-p / temperature sampling inside the decode loop (sampler.cu): scratch, the uniform of the current draw, and
    // {seed, draw counter} on the device; the graphs bake k/T/p, so they are re-captured when those change
    void *sample_scratch = nullptr;
    float *d_u = nullptr;
    uint64_t *d_draw = nullptr;
    b2_sampler_cfg graph_sampler{};
    // batched prompt evaluation (prefill.cu): activations of up to PF_MAX_TOKENS prompt tokens and the GEMM's B operand
    float *pf_x = nullptr, *pf_qkv = nullptr, *pf_attn = nullptr, *pf_h = nullptr;
    void *pf_pb = nullptr;
    float *pf_part = nullptr;      // split-K partial sums
    uint64_t pf_part_floats = 0;
    bool use_pf_gemm = true;
    // attention variants: [..][0] positions below ATTN_SHORT_CTX (one 128-thread CTA per head), [..][1] below ATTN_CTA_CTX (one 1024-thread
    // CTA per head, its eight groups split the context), [..][2] longer contexts (a cluster of eight CTAs per head)
    cudaGraphExec_t graphs[G_COUNT][ATTN_VARIANTS] = {};
    int graph_launches[G_COUNT][ATTN_VARIANTS] = {};
    // greedy decode also exists as a graph of `multi_n` consecutive tokens (B2_GRAPH_TOKENS, default 8): fewer graph launches,
    // no gap between the tokens of one graph
    cudaGraphExec_t multi[2][ATTN_VARIANTS] = {};   // [G_DECODE / G_DECODE_VP][attention variant]
    int multi_n = 8;
    bool use_graphs = true;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    b2_stats stats{};
    int host_pos = 0;
    bool have_logits = false;
    cudaError_t enqueue_error = cudaSuccess;   // first launch error of the tensor-parallel pass (reported by run_graph)
    std::vector<void *> allocs;
};

static int dev_alloc(std::vector<void *> &owner, void **p, uint64_t n) {
    cudaError_t e = cudaMalloc(p, n ? n : 16);
    if (e != cudaSuccess) {
        b2_set_error("cudaMalloc(%llu) failed: %s", (unsigned long long)n, cudaGetErrorString(e));
        return e == cudaErrorMemoryAllocation ? B2_ENOMEM : B2_ECUDA;
    }
    owner.push_back(*p);
    return B2_OK;
}

// ------------------------------------------------------------------------------------ engine
extern "C" int b2_engine_create(int device_ordinal, int tp_rank, int tp_size, b2_engine **out) {
    ApiScope _scope("b2_engine_create");
    B2_CHECK(out, B2_EINVAL, "b2_engine_create: out is NULL");
    B2_CHECK(tp_size == 1 || tp_size == 2 || tp_size == 4 || tp_size == 8, B2_EINVAL, "b2_engine_create: tensor parallel size %d (1, 2, 4 or 8)", tp_size);
    B2_CHECK(tp_rank >= 0 && tp_rank < tp_size, B2_EINVAL, "b2_engine_create: rank %d of %d", tp_rank, tp_size);
    int n = 0;
    B2_CUDA(cudaGetDeviceCount(&n));
    B2_CHECK(device_ordinal >= 0 && device_ordinal < n, B2_EINVAL, "b2_engine_create: device %d of %d", device_ordinal, n);
    B2_CUDA(cudaSetDevice(device_ordinal));
    cudaDeviceProp prop;
    B2_CUDA(cudaGetDeviceProperties(&prop, device_ordinal));
    B2_CHECK(prop.major == 10, B2_EINVAL, "b2_engine_create: device %d is sm_%d%d; this library is built for sm_100a only", device_ordinal, prop.major, prop.minor);
    b2_engine *e = new b2_engine();
    e->device = device_ordinal;
    e->sm_count = prop.multiProcessorCount;
    e->tp_rank = tp_rank;
    e->tp_size = tp_size;
    e->tp.n = tp_size;
    e->tp.rank = tp_rank;
    B2_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    *out = e;
    return B2_OK;
}
// The exchange block of this rank (allocated and zeroed on the first call) as a CUDA IPC handle for the other ranks' processes.
extern "C" int b2_engine_tp_handle(b2_engine *e, void *handle_out) {
    ApiScope _scope("b2_engine_tp_handle");
    B2_CHECK(e && handle_out, B2_EINVAL, "b2_engine_tp_handle: NULL");
    B2_CHECK(e->tp_size > 1, B2_EINVAL, "b2_engine_tp_handle: the engine was created with tp_size 1");
    static_assert(sizeof(cudaIpcMemHandle_t) == B2_TP_HANDLE_BYTES, "handle size");
    B2_CUDA(cudaSetDevice(e->device));
    if (!e->tp.local) {
        B2_CUDA(cudaMalloc(&e->tp.local, TpLink::bytes()));
        B2_CUDA(cudaMemset(e->tp.local, 0, TpLink::bytes()));
        B2_CUDA(cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t h;
    B2_CUDA(cudaIpcGetMemHandle(&h, e->tp.local));
    memcpy(handle_out, &h, sizeof h);
    return B2_OK;
}
// handles: tp_size consecutive handles, entry r from rank r's b2_engine_tp_handle (this rank's own entry is not opened).
extern "C" int b2_engine_tp_connect(b2_engine *e, const void *handles) {
    ApiScope _scope("b2_engine_tp_connect");
    B2_CHECK(e && handles, B2_EINVAL, "b2_engine_tp_connect: NULL");
    B2_CHECK(e->tp_size > 1 && e->tp.local, B2_EINVAL, "b2_engine_tp_connect: call b2_engine_tp_handle first");
    B2_CHECK(!e->tp.connected, B2_EINVAL, "b2_engine_tp_connect: already connected");
    B2_CUDA(cudaSetDevice(e->device));
    for (int r = 0; r < e->tp_size; ++r) {
        if (r == e->tp_rank) { e->tp.peer[r] = e->tp.local; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char *)handles + (size_t)r * sizeof h, sizeof h);
        cudaError_t err = cudaIpcOpenMemHandle(&e->tp.peer[r], h, cudaIpcMemLazyEnablePeerAccess);
        if (err != cudaSuccess) {
            cudaGetLastError();
            b2_set_error("b2_engine_tp_connect: cannot map the exchange block of rank %d (%s); tensor parallelism needs peer access between the GPUs", r,
                         cudaGetErrorString(err));
            return B2_ECUDA;
        }
    }
    e->tp.connected = true;
    return B2_OK;
}
extern "C" int b2_engine_destroy(b2_engine *e) {
    ApiScope _scope("b2_engine_destroy");
    if (!e) return B2_OK;
    cudaSetDevice(e->device);
    if (e->tp.connected)
        for (int r = 0; r < e->tp_size; ++r)
            if (r != e->tp_rank && e->tp.peer[r]) cudaIpcCloseMemHandle(e->tp.peer[r]);
    if (e->tp.local) cudaFree(e->tp.local);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
    return B2_OK;
}
extern "C" int b2_engine_device(const b2_engine *e, int *d) {
    B2_CHECK(e && d, B2_EINVAL, "b2_engine_device: NULL");
    *d = e->device;
    return B2_OK;
}

// ------------------------------------------------------------------------------------ model
static int check_cfg(const b2_arch_cfg *c) {
    B2_CHECK(c, B2_EINVAL, "arch cfg is NULL");
    B2_CHECK(c->hidden > 0 && c->hidden % QK == 0, B2_EINVAL, "hidden %d must be a positive multiple of 32", c->hidden);
    B2_CHECK(c->ffn > 0 && c->ffn % QK == 0, B2_EINVAL, "ffn %d must be a positive multiple of 32", c->ffn);
    B2_CHECK(c->head_dim >= 32 && c->head_dim <= 256 && c->head_dim % 32 == 0, B2_EINVAL, "head_dim %d must be a multiple of 32 in [32,256]", c->head_dim);
    B2_CHECK(c->n_heads > 0 && c->n_kv_heads > 0 && c->n_heads % c->n_kv_heads == 0, B2_EINVAL, "heads %d / kv heads %d", c->n_heads, c->n_kv_heads);
    B2_CHECK(c->n_layers > 0 && c->vocab > 0, B2_EINVAL, "layers %d vocab %d", c->n_layers, c->vocab);
    B2_CHECK(c->max_ctx > 0, B2_EINVAL, "max_ctx %d", c->max_ctx);
    B2_CHECK(c->hidden <= 8192, B2_EINVAL, "hidden %d too large (the fused RMSNorm prologue stages at most 8192)", c->hidden);
    return B2_OK;
}

static int build_rope(b2_model *m) {
    const b2_arch_cfg &c = m->cfg;
    int half = c.head_dim / 2;
    std::vector<double> inv(half);
    for (int j = 0; j < half; ++j) {
        double f = 1.0 / pow((double)c.rope_theta, (2.0 * j) / c.head_dim);
        if (c.rope_llama3) {   // Llama-3.1 frequency scaling (oracle/model.py:rope_inv_freq)
            double factor = c.rope_factor, lo = c.rope_low_freq, hi = c.rope_high_freq, orig = c.rope_orig_ctx;
            double wavelen = 2.0 * M_PI / f, lo_wl = orig / lo, hi_wl = orig / hi;
            double scaled = wavelen > lo_wl ? f / factor : f;
            double smooth = (orig / wavelen - lo) / (hi - lo);
            double mid = (1.0 - smooth) * f / factor + smooth * f;
            bool is_mid = !(wavelen < hi_wl) && !(wavelen > lo_wl);
            f = is_mid ? mid : scaled;
        }
        inv[j] = f;
    }
    std::vector<float> cs((size_t)c.max_ctx * half), sn((size_t)c.max_ctx * half);
    for (int p = 0; p < c.max_ctx; ++p)
        for (int j = 0; j < half; ++j) {
            double a = (double)p * inv[j];
            cs[(size_t)p * half + j] = (float)cos(a);
            sn[(size_t)p * half + j] = (float)sin(a);
        }
    B2_TRY(dev_alloc(m->allocs, (void **)&m->rope_cos, cs.size() * 4));
    B2_TRY(dev_alloc(m->allocs, (void **)&m->rope_sin, sn.size() * 4));
    B2_CUDA(cudaMemcpy(m->rope_cos, cs.data(), cs.size() * 4, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(m->rope_sin, sn.data(), sn.size() * 4, cudaMemcpyHostToDevice));
    return B2_OK;
}

static inline uint64_t packed_bytes_of(int type, int64_t rows, int64_t cols) {
    return type == B2_T_Q6_K ? packed_q6_k_bytes(rows, cols) : type == B2_T_Q4_K ? packed_q4_k_bytes(rows, cols) : packed_q4_0_bytes(rows, cols);
}
static int alloc_mat(b2_model *m, Mat &mat, int64_t rows, int64_t cols, int type) {
    B2_CHECK(type == B2_T_Q4_0 || type == B2_T_Q4_K || type == B2_T_Q6_K, B2_EINVAL, "tensor type %d not supported (Q4_0, Q4_K and Q6_K matrices in this build)", type);
    B2_CHECK(type == B2_T_Q4_0 || cols % 256 == 0, B2_EINVAL, "K-quants need a multiple of 256 columns, got %lld", (long long)cols);
    mat.rows = rows;
    mat.cols = cols;
    mat.type = type;
    B2_TRY(dev_alloc(m->allocs, &mat.packed, packed_bytes_of(type, rows, cols)));
    B2_CUDA(cudaMemsetAsync(mat.packed, 0, packed_bytes_of(type, rows, cols), m->e->stream));   // row / block padding must be zero
    return B2_OK;
}
// GGML bytes of the matrix (what a token has to read): Q4_0 18 B / 32 weights, Q4_K 144 B / 256 weights -- both 4.5 bit
static inline uint64_t algo_bytes(const Mat &m) {
    return m.type == B2_T_Q6_K ? (uint64_t)m.rows * (m.cols / 256) * 210 : (uint64_t)m.rows * (m.cols / QK) * 18;
}

static int alloc_model_skeleton(b2_engine *e, const b2_arch_cfg *cfg, const ModelTypes &mt, b2_model **out) {
    B2_CHECK(e && out, B2_EINVAL, "NULL engine/out");
    B2_TRY(check_cfg(cfg));
    B2_CUDA(cudaSetDevice(e->device));
    const int tp = e->tp_size;
    if (tp > 1) {
        B2_CHECK(e->tp.connected, B2_EINVAL, "tensor parallel engine: call b2_engine_tp_connect before creating a model");
        B2_CHECK(cfg->n_kv_heads % tp == 0, B2_EINVAL, "%d KV heads do not split over %d ranks", cfg->n_kv_heads, tp);
        B2_CHECK(cfg->ffn % (QK * tp) == 0, B2_EINVAL, "ffn %d does not split over %d ranks in whole blocks", cfg->ffn, tp);
        B2_CHECK(cfg->hidden <= TP_MAX_VEC, B2_EINVAL, "hidden %d exceeds the exchange slots (%d)", cfg->hidden, TP_MAX_VEC);
        bool kq = false;
        for (const LayerTypes &lt : mt.layers) kq = kq || lt.o != B2_T_Q4_0 || lt.down != B2_T_Q4_0;
        if (kq) {   // a column window of a K-quant matrix must start and end on a 256-element super-block
            B2_CHECK((cfg->n_heads / tp * cfg->head_dim) % 256 == 0 && (cfg->ffn / tp) % 256 == 0, B2_EINVAL,
                     "K-quant O / down matrices do not split over %d ranks in whole super-blocks", tp);
        }
    }
    b2_model *m = new b2_model();
    m->e = e;
    m->cfg = *cfg;
    m->tp = tp;
    m->tp_rank = e->tp_rank;
    const b2_arch_cfg &c = m->cfg;
    int H = c.hidden, I = m->ffn(), qd = m->q_dim(), kvd = m->kv_dim();
    int rc = B2_OK;
    do {
        if ((rc = alloc_mat(m, m->embd, c.vocab, H, mt.embd))) break;
        if (c.tied) m->head = m->embd;
        else if ((rc = alloc_mat(m, m->head, c.vocab, H, mt.head))) break;
        if ((rc = dev_alloc(m->allocs, (void **)&m->out_norm, H * 4))) break;
        m->layers.resize(c.n_layers);
        for (size_t li = 0; li < m->layers.size(); ++li) {
            LayerW &ly = m->layers[li];
            const LayerTypes &lt = mt.layers[li];
            if ((rc = dev_alloc(m->allocs, (void **)&ly.attn_norm, H * 4))) break;
            if ((rc = dev_alloc(m->allocs, (void **)&ly.ffn_norm, H * 4))) break;
            if (c.qkv_bias && (rc = dev_alloc(m->allocs, (void **)&ly.qkv_bias, (qd + 2 * kvd) * 4))) break;
            if (lt.q == lt.k && lt.k == lt.v) {
                if ((rc = alloc_mat(m, ly.qkv, qd + 2 * kvd, H, lt.q))) break;
            } else {
                ly.split_qkv = true;
                if ((rc = alloc_mat(m, ly.qkv_part[0], qd, H, lt.q))) break;
                if ((rc = alloc_mat(m, ly.qkv_part[1], kvd, H, lt.k))) break;
                if ((rc = alloc_mat(m, ly.qkv_part[2], kvd, H, lt.v))) break;
            }
            if ((rc = alloc_mat(m, ly.o, H, qd, lt.o))) break;
            if ((rc = alloc_mat(m, ly.gateup, 2 * (int64_t)I, H, lt.gateup))) break;
            if ((rc = alloc_mat(m, ly.down, H, I, lt.down))) break;
        }
        if (rc) break;
        if ((rc = build_rope(m))) break;
    } while (0);
    if (rc) { b2_model_destroy(m); return rc; }
    // algorithmic bytes per decoded token: every matmul weight once (4.5 bit/weight), LM head included
    uint64_t w = 0;
    for (auto &ly : m->layers)
        w += (ly.split_qkv ? algo_bytes(ly.qkv_part[0]) + algo_bytes(ly.qkv_part[1]) + algo_bytes(ly.qkv_part[2]) : algo_bytes(ly.qkv)) +
             algo_bytes(ly.o) + algo_bytes(ly.gateup) + algo_bytes(ly.down);
    w += algo_bytes(m->head);
    m->weight_bytes = w;
    *out = m;
    return B2_OK;
}

extern "C" int b2_model_destroy(b2_model *m) {
    ApiScope _scope("b2_model_destroy");
    if (!m) return B2_OK;
    if (m->e) cudaSetDevice(m->e->device);
    for (void *p : m->allocs) cudaFree(p);
    delete m;
    return B2_OK;
}
extern "C" int b2_model_weight_bytes(const b2_model *m, uint64_t *out) {
    B2_CHECK(m && out, B2_EINVAL, "NULL");
    *out = m->weight_bytes;
    return B2_OK;
}

// tensor ids of the random initialiser (oracle/archs.py:KIND)
enum { K_EMBD = 0, K_OUT_NORM, K_OUTPUT, K_ATTN_NORM, K_Q, K_K, K_V, K_QB, K_KB, K_VB, K_O, K_FFN_NORM, K_GATE, K_UP, K_DOWN };
#define GLOBAL_LAYER 0xFFFFu
static inline uint32_t tid_of(int kind, uint32_t layer) { return layer * 16 + kind; }

extern "C" int b2_model_create_random(b2_engine *e, const b2_arch_cfg *cfg, uint64_t seed, int qtype, int head_qtype, b2_model **out) {
    ApiScope _scope("b2_model_create_random");
    b2_model *m = nullptr;
    B2_TRY(alloc_model_skeleton(e, cfg, ModelTypes::uniform(cfg ? cfg->n_layers : 0, qtype, head_qtype), &m));
    const b2_arch_cfg &c = m->cfg;
    cudaStream_t st = e->stream;
    int H = c.hidden, I = m->ffn(), qd = m->q_dim(), kvd = m->kv_dim();
    float nsig = 0.1f, noff = c.norm_plus_one ? 0.0f : 1.0f, wsig = 0.02f, bsig = 0.1f;
    auto key = [&](int kind, uint32_t layer) { return ri_tensor_key(seed, tid_of(kind, layer)); };
    // Q4_0: gaussian weights through ggml's quantiser (bit-identical to b2_randinit_tensor and oracle/randinit.py);
    // Q4_K: synthetic blocks (gemv.cu:k_randinit_q4_k), for benchmarks only
    // (rows x cols) window at (src_row0, src_col0) of the whole tensor, whose rows are src_cols wide, into rows dst_row0, +stride, ...
    auto ri = [&](const Mat &mt, uint64_t k, float sig, int64_t rows, int64_t cols, int64_t row0, int64_t stride, int64_t src_row0 = 0,
                  int64_t src_cols = 0, int64_t src_col0 = 0) {
        if (mt.type == B2_T_Q6_K) launch_randinit_q6_k(k, rows, cols, mt.packed, row0, stride, st, src_row0, src_cols, src_col0);
        else if (mt.type == B2_T_Q4_K) launch_randinit_q4_k(k, rows, cols, mt.packed, row0, stride, st, src_row0, src_cols, src_col0);
        else launch_randinit_q4_0(k, sig, rows, cols, mt.packed, row0, stride, st, src_row0, src_cols, src_col0);
    };
    const int64_t rk = m->tp_rank, qd_full = (int64_t)c.n_heads * c.head_dim, I_full = c.ffn;
    ri(m->embd, key(K_EMBD, GLOBAL_LAYER), wsig, c.vocab, H, 0, 1);
    if (!c.tied) ri(m->head, key(K_OUTPUT, GLOBAL_LAYER), wsig, c.vocab, H, 0, 1);
    launch_randinit_f32(key(K_OUT_NORM, GLOBAL_LAYER), nsig, noff, H, m->out_norm, st);
    for (int l = 0; l < c.n_layers; ++l) {
        LayerW &ly = m->layers[l];
        launch_randinit_f32(key(K_ATTN_NORM, l), nsig, noff, H, ly.attn_norm, st);
        launch_randinit_f32(key(K_FFN_NORM, l), nsig, noff, H, ly.ffn_norm, st);
        ri(ly.qkv, key(K_Q, l), wsig, qd, H, 0, 1, rk * qd);
        ri(ly.qkv, key(K_K, l), wsig, kvd, H, qd, 1, rk * kvd);
        ri(ly.qkv, key(K_V, l), wsig, kvd, H, qd + kvd, 1, rk * kvd);
        if (c.qkv_bias) {
            launch_randinit_f32(key(K_QB, l), bsig, 0.f, qd, ly.qkv_bias, st, rk * qd);
            launch_randinit_f32(key(K_KB, l), bsig, 0.f, kvd, ly.qkv_bias + qd, st, rk * kvd);
            launch_randinit_f32(key(K_VB, l), bsig, 0.f, kvd, ly.qkv_bias + qd + kvd, st, rk * kvd);
        }
        ri(ly.o, key(K_O, l), wsig, H, qd, 0, 1, 0, qd_full, rk * qd);
        ri(ly.gateup, key(K_GATE, l), wsig, I, H, 0, 2, rk * I);
        ri(ly.gateup, key(K_UP, l), wsig, I, H, 1, 2, rk * I);
        ri(ly.down, key(K_DOWN, l), wsig, H, I, 0, 1, 0, I_full, rk * I);
    }
    cudaError_t err = cudaStreamSynchronize(st);
    if (err == cudaSuccess) err = cudaGetLastError();
    if (err != cudaSuccess) {
        b2_set_error("b2_model_create_random: %s", cudaGetErrorString(err));
        b2_model_destroy(m);
        return B2_ECUDA;
    }
    *out = m;
    return B2_OK;
}

struct Staging {
    void *buf = nullptr;
    uint64_t cap = 0;
    ~Staging() { if (buf) cudaFree(buf); }
    int ensure(uint64_t n) {
        if (n <= cap) return B2_OK;
        if (buf) cudaFree(buf);
        buf = nullptr;
        cap = 0;
        B2_CUDA(cudaMalloc(&buf, n));
        cap = n;
        return B2_OK;
    }
};

// rows [src_row0, src_row0 + nrows) of the tensor; src_col0 >= 0: only its columns [src_col0, src_col0 + dst.cols) (a rank's K slice)
static int upload_mat(b2_model *m, Staging &sg, const b2_tensor_desc &t, Mat &dst, int64_t dst_row0, int64_t stride, int64_t src_row0, int64_t nrows,
                      int64_t src_col0 = -1) {
    B2_CHECK(t.type == dst.type, B2_EINVAL, "tensor %s: type %d, but the matrix it belongs to is type %d (all parts of a fused matrix share one format)", t.name, t.type, dst.type);
    if (src_col0 < 0) B2_CHECK(t.cols == dst.cols, B2_EINVAL, "tensor %s: cols %lld, expected %lld", t.name, (long long)t.cols, (long long)dst.cols);
    else B2_CHECK(src_col0 + dst.cols <= t.cols, B2_EINVAL, "tensor %s: cols %lld, need %lld", t.name, (long long)t.cols, (long long)(src_col0 + dst.cols));
    uint64_t row_bytes = t.type == B2_T_Q6_K ? (uint64_t)(t.cols / 256) * 210 : t.type == B2_T_Q4_K ? (uint64_t)(t.cols / 256) * 144 : (uint64_t)(t.cols / QK) * 18;
    B2_CHECK(t.nbytes == (uint64_t)t.rows * row_bytes, B2_EINVAL, "tensor %s: %llu bytes, expected %llu", t.name, (unsigned long long)t.nbytes, (unsigned long long)(t.rows * row_bytes));
    B2_CHECK(src_row0 + nrows <= t.rows, B2_EINVAL, "tensor %s: has %lld rows, need %lld", t.name, (long long)t.rows, (long long)(src_row0 + nrows));
    B2_TRY(sg.ensure(nrows * row_bytes));
    B2_CUDA(cudaMemcpyAsync(sg.buf, (const char *)t.data + src_row0 * row_bytes, nrows * row_bytes, cudaMemcpyHostToDevice, m->e->stream));
    const int64_t wc = src_col0 < 0 ? 0 : t.cols, c0 = src_col0 < 0 ? 0 : src_col0;
    if (t.type == B2_T_Q6_K) launch_pack_q6_k(sg.buf, nrows, dst.cols, dst.packed, dst_row0, stride, m->e->stream, wc, c0);
    else if (t.type == B2_T_Q4_K) launch_pack_q4_k(sg.buf, nrows, dst.cols, dst.packed, dst_row0, stride, m->e->stream, wc, c0);
    else launch_pack_q4_0(sg.buf, nrows, dst.cols, dst.packed, dst_row0, stride, m->e->stream, wc, c0);
    B2_CUDA(cudaStreamSynchronize(m->e->stream));
    return B2_OK;
}
static int upload_vec(b2_model *m, const b2_tensor_desc &t, float *dst, int64_t n, int64_t src0 = 0, int64_t full = -1) {
    if (full < 0) full = n;
    B2_CHECK(t.type == B2_T_F32 && t.rows == full && t.nbytes == (uint64_t)full * 4, B2_EINVAL, "tensor %s: expected f32[%lld]", t.name, (long long)full);
    B2_CUDA(cudaMemcpy(dst, (const float *)t.data + src0, n * 4, cudaMemcpyHostToDevice));
    (void)m;
    return B2_OK;
}

extern "C" int b2_model_create(b2_engine *e, const b2_arch_cfg *cfg, const b2_tensor_desc *tensors, int n_tensors, b2_model **out) {
    ApiScope _scope("b2_model_create");
    B2_CHECK(tensors && n_tensors > 0, B2_EINVAL, "b2_model_create: no tensors");
    std::map<std::string, const b2_tensor_desc *> by_name;
    for (int i = 0; i < n_tensors; ++i) {
        B2_CHECK(tensors[i].name && tensors[i].data, B2_EINVAL, "tensor %d has NULL name/data", i);
        by_name[tensors[i].name] = &tensors[i];
    }
    auto find = [&](const std::string &n) -> const b2_tensor_desc * {
        auto it = by_name.find(n);
        return it == by_name.end() ? nullptr : it->second;
    };
    const b2_tensor_desc *te = find("token_embd.weight");
    B2_CHECK(te, B2_EINVAL, "missing tensor token_embd.weight");
    const b2_tensor_desc *l0 = find("blk.0.attn_q.weight");
    if (!l0) l0 = find("blk.0.attn_qkv.weight");
    B2_CHECK(l0, B2_EINVAL, "missing tensor blk.0.attn_q.weight / blk.0.attn_qkv.weight");
    b2_model *m = nullptr;
    B2_CHECK(cfg && cfg->n_layers > 0 && cfg->n_layers < 4096, B2_EINVAL, "b2_model_create: bad config");
    // every matrix keeps the format its tensor has in the file (llama.cpp mixes formats inside one model)
    ModelTypes mt = ModelTypes::uniform(cfg->n_layers, l0->type, te->type);
    if (const b2_tensor_desc *ow = find("output.weight")) mt.head = ow->type;
    for (int l = 0; l < cfg->n_layers; ++l) {
        const std::string p = "blk." + std::to_string(l) + ".";
        LayerTypes &lt = mt.layers[l];
        auto ty = [&](const char *n, int dflt) { const b2_tensor_desc *t = find(p + n); return t ? t->type : dflt; };
        if (const b2_tensor_desc *f = find(p + "attn_qkv.weight")) lt.q = lt.k = lt.v = f->type;
        else { lt.q = ty("attn_q.weight", lt.q); lt.k = ty("attn_k.weight", lt.q); lt.v = ty("attn_v.weight", lt.q); }
        lt.o = ty("attn_output.weight", lt.q);
        lt.gateup = ty("ffn_up.weight", lt.q);
        lt.down = ty("ffn_down.weight", lt.q);
    }
    B2_TRY(alloc_model_skeleton(e, cfg, mt, &m));
    const b2_arch_cfg &c = m->cfg;
    int H = c.hidden, I = m->ffn(), qd = m->q_dim(), kvd = m->kv_dim();
    Staging sg;
    int rc = B2_OK;
    // a tensor-parallel rank takes its rows of q / k / v / gate / up and its columns of attn_output / ffn_down
    const int64_t rk = m->tp_rank, qdF = (int64_t)c.n_heads * c.head_dim, kvdF = (int64_t)c.n_kv_heads * c.head_dim, IF = c.ffn;
    const int64_t ocol = m->tp > 1 ? rk * qd : -1, dcol = m->tp > 1 ? rk * I : -1;
#define NEED(var, name)                                                        \
    const b2_tensor_desc *var = find(name);                                    \
    if (!var) { b2_set_error("missing tensor %s", std::string(name).c_str()); rc = B2_EINVAL; break; }
    do {
        if ((rc = upload_mat(m, sg, *te, m->embd, 0, 1, 0, c.vocab))) break;
        NEED(on, "output_norm.weight");
        if ((rc = upload_vec(m, *on, m->out_norm, H))) break;
        if (!c.tied) {
            NEED(ow, "output.weight");
            if ((rc = upload_mat(m, sg, *ow, m->head, 0, 1, 0, c.vocab))) break;
        }
        for (int l = 0; l < c.n_layers && !rc; ++l) {
            LayerW &ly = m->layers[l];
            std::string p = "blk." + std::to_string(l) + ".";
            do {
                NEED(an, p + "attn_norm.weight");
                if ((rc = upload_vec(m, *an, ly.attn_norm, H))) break;
                NEED(fn, p + "ffn_norm.weight");
                if ((rc = upload_vec(m, *fn, ly.ffn_norm, H))) break;
                if (const b2_tensor_desc *f = find(p + "attn_qkv.weight")) {
                    if ((rc = (m->tp > 1 ? (upload_mat(m, sg, *f, ly.qkv, 0, 1, rk * qd, qd) || upload_mat(m, sg, *f, ly.qkv, qd, 1, qdF + rk * kvd, kvd) || upload_mat(m, sg, *f, ly.qkv, qd + kvd, 1, qdF + kvdF + rk * kvd, kvd)) : upload_mat(m, sg, *f, ly.qkv, 0, 1, 0, qd + 2 * kvd)))) break;
                } else {
                    NEED(tq, p + "attn_q.weight");
                    NEED(tk, p + "attn_k.weight");
                    NEED(tv, p + "attn_v.weight");
                    if (ly.split_qkv) {
                        if ((rc = upload_mat(m, sg, *tq, ly.qkv_part[0], 0, 1, rk * qd, qd))) break;
                        if ((rc = upload_mat(m, sg, *tk, ly.qkv_part[1], 0, 1, rk * kvd, kvd))) break;
                        if ((rc = upload_mat(m, sg, *tv, ly.qkv_part[2], 0, 1, rk * kvd, kvd))) break;
                    } else {
                        if ((rc = upload_mat(m, sg, *tq, ly.qkv, 0, 1, rk * qd, qd))) break;
                        if ((rc = upload_mat(m, sg, *tk, ly.qkv, qd, 1, rk * kvd, kvd))) break;
                        if ((rc = upload_mat(m, sg, *tv, ly.qkv, qd + kvd, 1, rk * kvd, kvd))) break;
                    }
                }
                if (c.qkv_bias) {
                    NEED(bq, p + "attn_q.bias");
                    NEED(bk, p + "attn_k.bias");
                    NEED(bv, p + "attn_v.bias");
                    if ((rc = upload_vec(m, *bq, ly.qkv_bias, qd, rk * qd, qdF))) break;
                    if ((rc = upload_vec(m, *bk, ly.qkv_bias + qd, kvd, rk * kvd, kvdF))) break;
                    if ((rc = upload_vec(m, *bv, ly.qkv_bias + qd + kvd, kvd, rk * kvd, kvdF))) break;
                }
                NEED(to, p + "attn_output.weight");
                if ((rc = upload_mat(m, sg, *to, ly.o, 0, 1, 0, H, ocol))) break;
                const b2_tensor_desc *tg = find(p + "ffn_gate.weight");
                NEED(tu, p + "ffn_up.weight");
                if (tg) {
                    if ((rc = upload_mat(m, sg, *tg, ly.gateup, 0, 2, rk * I, I))) break;
                    if ((rc = upload_mat(m, sg, *tu, ly.gateup, 1, 2, rk * I, I))) break;
                } else {   // phi3-style fused [gate; up]
                    if ((rc = upload_mat(m, sg, *tu, ly.gateup, 0, 2, rk * I, I))) break;
                    if ((rc = upload_mat(m, sg, *tu, ly.gateup, 1, 2, IF + rk * I, I))) break;
                }
                NEED(td, p + "ffn_down.weight");
                if ((rc = upload_mat(m, sg, *td, ly.down, 0, 1, 0, H, dcol))) break;
            } while (0);
        }
    } while (0);
#undef NEED
    if (rc) { b2_model_destroy(m); return rc; }
    *out = m;
    return B2_OK;
}

// ------------------------------------------------------------------------------------ session
// (these bookkeeping kernels are plain stream-ordered launches: they start after the previous kernel has completed)
__global__ void k_load_prompt(DevState *st, const int *prompt) {
    st->cur_token = prompt[st->prompt_idx];
    st->prompt_idx += 1;
}
__global__ void k_commit_sample(DevState *st, const int *sampled, int *out_tokens, int *ctx) {
    int t = *sampled;
    st->cur_token = t;
    out_tokens[st->step] = t;
    ctx[st->pos] = t;   // the forward pass that follows puts it at position pos
}
__global__ void k_advance(DevState *st, int inc_step) {
    st->pos += 1;
    st->step += inc_step;
}
// uniform of draw i of a session: top 24 bits of splitmix64(splitmix64(seed) + i)  (oracle/sampler.py:uniform)
__global__ void k_make_u(uint64_t *draw, float *u) {
    const uint64_t h = ri_splitmix64(ri_splitmix64(draw[0]) + draw[1]);
    *u = (float)(h >> 40) * 5.9604644775390625e-08f;
    draw[1] += 1;
}
__global__ void k_copy_logits(const float *logits, int n, float *log, const DevState *st) {
    float *dst = log + (int64_t)st->step * n;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = logits[i];
}

// CTAs (= arg-max partials) of the LM-head GEMV of a tensor-parallel rank: replicated over the whole vocabulary, or over its row window
static int tp_head_grid(const b2_model *m, bool vp) { return gemv_grid((int)ntiles(vp ? m->cfg.vocab / m->tp : m->cfg.vocab), 1); }
static bool tp_vocab_parallel(const b2_model *m) {
    static int off = -1;   // B2_TP_REPLICATED_HEAD=1: every rank multiplies the whole LM head in every step
    if (off < 0) { const char *e = getenv("B2_TP_REPLICATED_HEAD"); off = (e && e[0] == '1') ? 1 : 0; }
    return m->tp > 1 && !off && m->cfg.vocab % (TILE_ROWS * m->tp) == 0;
}
// greedy token of the step: arg-max over the partials of the last LM-head GEMV (tensor parallel: exchanged between the ranks)
static void enqueue_greedy_pick(b2_session *s) {
    b2_model *m = s->m;
    cudaStream_t st = m->e->stream;
    if (m->tp == 1) { launch_argmax_partials(s->pmax, s->pidx, s->head_grid, s->d_sampled, st); return; }
    const TpLink &L = m->e->tp;
    void *peers[TP_MAX_RANKS] = {};
    for (int p = 0; p < L.n; ++p) peers[p] = L.amax(p);
    launch_tp_argmax(s->pmax, s->pidx, std::max(tp_head_grid(m, false), tp_head_grid(m, true)), peers, L.n, L.rank, L.amax(L.rank), L.forwards(), L.error(),
                     s->d_sampled, st);
}

// The same pass on one rank of a tensor-parallel group.  Per layer: QKV of this rank's heads, attention over them, O projection
// of this rank's columns (partial rows -> every rank's slots, buffer 0), gate/up of this rank's rows (its prologue forms
// x = residual + sum of the slots), down projection of this rank's columns (-> buffer 1), consumed by the next layer's QKV or by
// the LM head (replicated: every rank computes all logits from the same x, so all ranks sample the same token without a further
// exchange).  The residual alternates between two buffers: a consumer's CTA 0 writes the new one while the others read the old.
static int enqueue_forward_tp(b2_session *s, bool with_head, int attn_ctx, bool vp) {
    b2_model *m = s->m;
    const b2_arch_cfg &c = m->cfg;
    const TpLink &L = m->e->tp;
    cudaStream_t st = m->e->stream;
    const bool pdl = s->use_pdl;
    const int H = c.hidden, I = m->ffn(), qd = m->q_dim(), hd = c.head_dim, N = L.n, rk = L.rank;
    int n = 0;
    float *X[2] = {s->x, s->x + H};
    int cur = 0;
    const float escale = c.embed_scale ? sqrtf((float)H) : 1.0f;
    launch_embed_row(qt_of(m->embd.type), m->embd.packed, m->embd.cols, &s->state->cur_token, escale, X[0], s->xsq, st); n++;
    const uint64_t layer_kv = (uint64_t)s->n_pages * m->kv_heads() * PAGE_TOKENS * hd;
    const int rounds = 2 * c.n_layers;   // exchange rounds of a forward pass: O and down projection of every layer
    auto producer = [&](int buf, int round) {
        GemvTp t;
        t.n = N;
        for (int p = 0; p < N; ++p) t.dst[p] = L.slot(p, buf, rk);
        t.fwd = L.forwards(); t.round = round; t.rounds = rounds;
        return t;
    };
    auto consumer = [&](int buf, int round) {
        GemvTp t;
        t.n = N;
        t.res = X[cur]; t.xout = X[cur ^ 1];
        t.slots = L.slot(rk, buf, 0); t.slot_stride = TP_MAX_VEC;
        t.fwd = L.forwards();
        t.round = round; t.rounds = rounds;
        t.err = L.error();
        cur ^= 1;
        return t;
    };
    cudaError_t e = cudaSuccess;
    for (int l = 0; l < c.n_layers && e == cudaSuccess; ++l) {
        LayerW &ly = m->layers[l];
        if (l == 0)
            e = launch_gemv_qt(qt_of(ly.qkv.type), GEMV_IN_NORM, GEMV_OUT_BIAS, ly.qkv.packed, ly.qkv.rows, H, X[0], ly.attn_norm, c.rms_eps, c.norm_plus_one, nullptr,
                               s->qkv, ly.qkv_bias, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 0, s->xsq, 1, nullptr, st, pdl);
        else
            e = launch_gemv_tp(consumer(1, 2 * l - 1), qt_of(ly.qkv.type), GEMV_IN_TPNORM, GEMV_OUT_BIAS, ly.qkv.packed, ly.qkv.rows, H, ly.attn_norm, c.rms_eps,
                               c.norm_plus_one, nullptr, s->qkv, ly.qkv_bias, 0, nullptr, nullptr, nullptr, st, pdl);
        n++;
        if (e != cudaSuccess) break;
        e = launch_attn_fused(s->qkv, m->heads(), m->kv_heads(), hd, &s->state->pos, 0, m->rope_cos, m->rope_sin, s->kcache + l * layer_kv,
                              s->vcache + l * layer_kv, s->page_table, nullptr, s->act_attn, attn_ctx, st, pdl); n++;
        if (e != cudaSuccess) break;
        e = launch_gemv_tp(producer(0, 2 * l), qt_of(ly.o.type), GEMV_IN_ACT, GEMV_OUT_TPPART, ly.o.packed, H, qd, nullptr, 0.f, 0, s->act_attn, nullptr, nullptr, 0,
                           nullptr, nullptr, nullptr, st, pdl); n++;
        if (e != cudaSuccess) break;
        e = launch_gemv_tp(consumer(0, 2 * l), qt_of(ly.gateup.type), GEMV_IN_TPNORM, GEMV_OUT_ACTMUL, ly.gateup.packed, 2 * (int64_t)I, H, ly.ffn_norm, c.rms_eps,
                           c.norm_plus_one, nullptr, s->h, nullptr, c.act, nullptr, nullptr, nullptr, st, pdl); n++;
        if (e != cudaSuccess) break;
        e = launch_gemv_tp(producer(1, 2 * l + 1), qt_of(ly.down.type), GEMV_IN_F32, GEMV_OUT_TPPART, ly.down.packed, H, I, nullptr, 0.f, 0, s->h, nullptr, nullptr, 0,
                           nullptr, nullptr, nullptr, st, pdl); n++;
    }
    if (e == cudaSuccess && with_head && vp) {
        // vocabulary-parallel LM head: this rank's window of rows (whole 16-row tiles, so the window is a pointer offset into the
        // packed matrix); logits land in their place of the full vector, arg-max partials carry indices of the whole vocabulary
        const int64_t Vl = c.vocab / N, row0 = (int64_t)rk * Vl;
        const uint64_t off = (uint64_t)(row0 / TILE_ROWS) * nblocks_padded(H) * (packed_bytes_of(m->head.type, TILE_ROWS, H) / nblocks_padded(H));
        int gvp = 0;
        if (tp_head_grid(m, true) < tp_head_grid(m, false)) { launch_fill_partials(s->pmax, s->pidx, 1024, st); n++; }   // tiny vocabularies: fewer CTAs than the replicated head left partials for
        e = launch_gemv_tp(consumer(1, 2 * c.n_layers - 1), qt_of(m->head.type), GEMV_IN_TPNORM, GEMV_OUT_LOGITS, (const char *)m->head.packed + off, Vl, H,
                           m->out_norm, c.rms_eps, c.norm_plus_one, nullptr, s->logits + row0, nullptr, 0, s->pmax, s->pidx, &gvp, st, pdl, (int)row0); n++;
    } else if (e == cudaSuccess && with_head) {
        e = launch_gemv_tp(consumer(1, 2 * c.n_layers - 1), qt_of(m->head.type), GEMV_IN_TPNORM, GEMV_OUT_LOGITS, m->head.packed, c.vocab, H, m->out_norm, c.rms_eps,
                           c.norm_plus_one, nullptr, s->logits, nullptr, 0, s->pmax, s->pidx, &s->head_grid, st, pdl); n++;
    }
    if (e != cudaSuccess) { s->enqueue_error = e; return n; }
    // a pass without the LM head still has to see the last round complete before the next pass reuses the slots
    launch_tp_close(L.slot(rk, 1, 0), TP_MAX_VEC, N, H, L.forwards(), rounds, 2 * c.n_layers - 1, !with_head, L.error(), st); n++;
    return n;
}

// enqueue one forward pass for the token in state->cur_token at position state->pos; returns #launches
static int enqueue_forward(b2_session *s, bool with_head, int attn_ctx, bool vp = false) {
    if (s->m->tp > 1) return enqueue_forward_tp(s, with_head, attn_ctx, vp);
    b2_model *m = s->m;
    const b2_arch_cfg &c = m->cfg;
    cudaStream_t st = m->e->stream;
    const bool pdl = s->use_pdl;
    int H = c.hidden, I = m->ffn(), qd = m->q_dim(), hd = c.head_dim;
    int n = 0;
    auto bytes_of = [](const Mat &mt) { return packed_bytes_of(mt.type, mt.rows, mt.cols); };
    float escale = c.embed_scale ? sqrtf((float)H) : 1.0f;
    launch_embed_row(qt_of(m->embd.type), m->embd.packed, m->embd.cols, &s->state->cur_token, escale, s->x, s->xsq, st); n++;
    int xsq_n = 1;   // how many partial sums of squares the producer of the current x left in s->xsq
    uint64_t layer_kv = (uint64_t)s->n_pages * m->kv_heads() * PAGE_TOKENS * hd;
    for (int l = 0; l < c.n_layers; ++l) {
        LayerW &ly = m->layers[l];
        int g = 0;
        // each GEMV hints the matrix the next GEMV will stream (L2 prefetch of the head of every CTA's share)
        const Mat *after_down = (l + 1 < c.n_layers) ? &m->layers[l + 1].qkv : (with_head ? &m->head : nullptr);
        if (ly.split_qkv) {   // q, k, v in different formats: three GEMVs into the same QKV vector
            int64_t row0 = 0;
            for (int part = 0; part < 3; ++part) {
                const Mat &pm = ly.qkv_part[part];
                launch_gemv_qt(qt_of(pm.type), GEMV_IN_NORM, GEMV_OUT_BIAS, pm.packed, pm.rows, H, s->x, ly.attn_norm, c.rms_eps, c.norm_plus_one, nullptr,
                               s->qkv + row0, ly.qkv_bias ? ly.qkv_bias + row0 : nullptr, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 0, s->xsq, xsq_n,
                               nullptr, st, pdl); n++;
                row0 += pm.rows;
            }
        } else {
                    launch_gemv_qt(qt_of(ly.qkv.type), GEMV_IN_NORM, GEMV_OUT_BIAS, ly.qkv.packed, ly.qkv.rows, H, s->x, ly.attn_norm, c.rms_eps, c.norm_plus_one, nullptr,
                        s->qkv, ly.qkv_bias, nullptr, 0, nullptr, nullptr, nullptr, ly.o.packed, bytes_of(ly.o), s->xsq, xsq_n, nullptr, st, pdl); n++;
        }
        if (s->fused_attn) {
            launch_attn_fused(s->qkv, m->heads(), m->kv_heads(), hd, &s->state->pos, 0, m->rope_cos, m->rope_sin, s->kcache + l * layer_kv,
                              s->vcache + l * layer_kv, s->page_table, nullptr, s->act_attn, attn_ctx, st, pdl); n++;
        } else {
            launch_rope_kv(s->qkv, m->heads(), m->kv_heads(), hd, &s->state->pos, 0, m->rope_cos, m->rope_sin,
                           s->kcache + l * layer_kv, s->vcache + l * layer_kv, s->page_table, st); n++;
            launch_attn_decode(s->qkv, m->heads(), m->kv_heads(), hd, &s->state->pos, 0, s->kcache + l * layer_kv, s->vcache + l * layer_kv,
                               s->page_table, s->attn_out, s->attn_scratch, s->max_ctx, st); n++;
            launch_quant_act(s->attn_out, qd, s->act_attn, st); n++;
        }
        // two partial-sum buffers alternate: a kernel reads the partials of its input while writing those of its output
        launch_gemv_qt(qt_of(ly.o.type), GEMV_IN_ACT, GEMV_OUT_RES, ly.o.packed, H, qd, nullptr, nullptr, 0.f, 0, s->act_attn, s->x, s->x, nullptr, 0,
                    nullptr, nullptr, &g, ly.gateup.packed, bytes_of(ly.gateup), nullptr, 0, s->xsq + 512, st, pdl); n++;
        launch_gemv_qt(qt_of(ly.gateup.type), GEMV_IN_NORM, GEMV_OUT_ACTMUL, ly.gateup.packed, 2 * (int64_t)I, H, s->x, ly.ffn_norm, c.rms_eps, c.norm_plus_one, nullptr,
                    s->h, nullptr, nullptr, c.act, nullptr, nullptr, nullptr, ly.down.packed, bytes_of(ly.down), s->xsq + 512, g, nullptr, st, pdl); n++;
        launch_gemv_qt(qt_of(ly.down.type), GEMV_IN_F32, GEMV_OUT_RES, ly.down.packed, H, I, nullptr, nullptr, 0.f, 0, s->h, s->x, s->x, nullptr, 0,
                    nullptr, nullptr, &g, after_down ? after_down->packed : nullptr, after_down ? bytes_of(*after_down) : 0, nullptr, 0, s->xsq, st, pdl); n++;
        xsq_n = g;
    }
    if (with_head) {
        launch_gemv_qt(qt_of(m->head.type), GEMV_IN_NORM, GEMV_OUT_LOGITS, m->head.packed, c.vocab, H, s->x, m->out_norm, c.rms_eps, c.norm_plus_one, nullptr,
                    s->logits, nullptr, nullptr, 0, s->pmax, s->pidx, &s->head_grid, m->layers[0].qkv.packed, bytes_of(m->layers[0].qkv),
                    s->xsq, xsq_n, nullptr, st, pdl); n++;
    }
    return n;
}

// which attention kernel serves a position (attn.cu picks the kernel from the context bound it is launched with)
static inline int attn_variant(int pos) { return pos < ATTN_SHORT_CTX ? 0 : pos < ATTN_CTA_CTX ? 1 : 2; }
static int enqueue_graph_body(b2_session *s, int which, int variant) {
    const int bound = variant == 0 ? ATTN_SHORT_CTX : variant == 1 ? ATTN_CTA_CTX : s->max_ctx;   // the context the variant's attention kernel is sized for
    const int attn_ctx = s->max_ctx < bound ? s->max_ctx : bound;
    cudaStream_t st = s->m->e->stream;
    int n = 0;
    switch (which) {
    case G_PREFILL_BODY:
    case G_PREFILL_LAST:
    case G_FORCED:
        k_load_prompt<<<1, 1, 0, st>>>(s->state, s->d_prompt); n++;
        n += enqueue_forward(s, which != G_PREFILL_BODY, attn_ctx);
        if (which == G_FORCED) { k_copy_logits<<<64, 256, 0, st>>>(s->logits, s->m->cfg.vocab, s->logits_log, s->state); n++; }
        k_advance<<<1, 1, 0, st>>>(s->state, which == G_FORCED ? 1 : 0); n++;
        break;
    case G_SAMPLE:
    case G_SAMPLE_LOG:
        if (which == G_SAMPLE_LOG) { k_copy_logits<<<64, 256, 0, st>>>(s->logits, s->m->cfg.vocab, s->logits_log, s->state); n++; }
        k_make_u<<<1, 1, 0, st>>>(s->d_draw, s->d_u); n++;
        launch_sample_topk(s->logits, s->m->cfg.vocab, s->graph_sampler.top_k, s->graph_sampler.temperature, s->graph_sampler.top_p, s->d_u, 0.f,
                           s->d_sampled, s->sample_scratch, st, s->d_ctx, &s->state->pos, 0, s->graph_sampler.repeat_penalty,
                           s->graph_sampler.repeat_last_n); n++;
        k_commit_sample<<<1, 1, 0, st>>>(s->state, s->d_sampled, s->d_out_tokens, s->d_ctx); n++;
        launch_fill_partials(s->pmax, s->pidx, 1024, st); n++;   // a later greedy step reads every partial slot
        n += enqueue_forward(s, true, attn_ctx);
        k_advance<<<1, 1, 0, st>>>(s->state, 1); n++;
        break;
    case G_DECODE:
    case G_DECODE_LOG:
    case G_DECODE_VP:
        if (which == G_DECODE_LOG) { k_copy_logits<<<64, 256, 0, st>>>(s->logits, s->m->cfg.vocab, s->logits_log, s->state); n++; }
        enqueue_greedy_pick(s); n++;
        k_commit_sample<<<1, 1, 0, st>>>(s->state, s->d_sampled, s->d_out_tokens, s->d_ctx); n++;
        n += enqueue_forward(s, true, attn_ctx, which == G_DECODE_VP);
        k_advance<<<1, 1, 0, st>>>(s->state, 1); n++;
        break;
    }
    return n;
}

static void drop_graphs(b2_session *s) {
    for (int g = 0; g < G_COUNT; ++g)
        for (int v = 0; v < ATTN_VARIANTS; ++v)
            if (s->graphs[g][v]) { cudaGraphExecDestroy(s->graphs[g][v]); s->graphs[g][v] = nullptr; }
    for (int k = 0; k < 2; ++k)
        for (int v = 0; v < ATTN_VARIANTS; ++v)
            if (s->multi[k][v]) { cudaGraphExecDestroy(s->multi[k][v]); s->multi[k][v] = nullptr; }
}

// run graph `which` for `times` consecutive positions starting at `pos0` (captured lazily; plain launches when B2_NO_GRAPH=1).
// Positions below ATTN_SHORT_CTX use the variant whose attention runs one CTA per head.
static int run_graph(b2_session *s, int which, int times, int pos0) {
    cudaStream_t st = s->m->e->stream;
    for (int i = 0; i < times; ++i) {
        const int v = attn_variant(pos0 + i);
        // several greedy tokens in one graph while they all use the same attention variant
        const int U = s->multi_n;
        if ((which == G_DECODE || which == G_DECODE_VP) && s->use_graphs && U > 1 && times - i >= U && attn_variant(pos0 + i + U - 1) == v) {
            const int mk = which == G_DECODE_VP ? 1 : 0;
            if (!s->multi[mk][v]) {
                cudaGraph_t g = nullptr;
                B2_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
                int nl = 0;
                for (int u = 0; u < U; ++u) nl += enqueue_graph_body(s, which, v);
                s->graph_launches[which][v] = nl / U;
                cudaError_t e1 = cudaStreamEndCapture(st, &g);
                if (e1 == cudaSuccess && s->enqueue_error != cudaSuccess) { e1 = s->enqueue_error; cudaGraphDestroy(g); }
                if (e1 != cudaSuccess) { b2_set_error("graph capture failed: %s", cudaGetErrorString(e1)); return B2_ECUDA; }
                cudaError_t e2 = cudaGraphInstantiate(&s->multi[mk][v], g, 0);
                cudaGraphDestroy(g);
                if (e2 != cudaSuccess) { b2_set_error("graph instantiate failed: %s", cudaGetErrorString(e2)); return B2_ECUDA; }
            }
            B2_CUDA(cudaGraphLaunch(s->multi[mk][v], st));
            s->stats.kernel_launches += (uint64_t)s->graph_launches[which][v] * U;
            i += U - 1;
            continue;
        }
        if (!s->use_graphs) {
            s->graph_launches[which][v] = enqueue_graph_body(s, which, v);
            B2_CUDA(cudaGetLastError());
            B2_CUDA(s->enqueue_error);
        } else {
            if (!s->graphs[which][v]) {
                cudaGraph_t g = nullptr;
                B2_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
                s->graph_launches[which][v] = enqueue_graph_body(s, which, v);
                cudaError_t e1 = cudaStreamEndCapture(st, &g);
                if (e1 == cudaSuccess && s->enqueue_error != cudaSuccess) { e1 = s->enqueue_error; cudaGraphDestroy(g); }
                if (e1 != cudaSuccess) { b2_set_error("graph capture failed: %s", cudaGetErrorString(e1)); return B2_ECUDA; }
                cudaError_t e2 = cudaGraphInstantiate(&s->graphs[which][v], g, 0);
                cudaGraphDestroy(g);
                if (e2 != cudaSuccess) { b2_set_error("graph instantiate failed: %s", cudaGetErrorString(e2)); return B2_ECUDA; }
            }
            B2_CUDA(cudaGraphLaunch(s->graphs[which][v], st));
        }
        s->stats.kernel_launches += (uint64_t)s->graph_launches[which][v];
    }
    return B2_OK;
}

extern "C" int b2_session_create(b2_model *m, int max_ctx, b2_session **out) {
    ApiScope _scope("b2_session_create");
    B2_CHECK(m && out, B2_EINVAL, "b2_session_create: NULL");
    B2_CHECK(max_ctx > 0 && max_ctx <= m->cfg.max_ctx, B2_EINVAL, "max_ctx %d out of range (model max %d)", max_ctx, m->cfg.max_ctx);
    B2_CUDA(cudaSetDevice(m->e->device));
    const b2_arch_cfg &c = m->cfg;
    b2_session *s = new b2_session();
    s->m = m;
    s->max_ctx = max_ctx;
    s->n_pages = (max_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const char *ng = getenv("B2_NO_GRAPH");
    s->use_graphs = !(ng && ng[0] == '1');
    const char *np = getenv("B2_NO_PDL");
    s->use_pdl = !(np && np[0] == '1');
    const char *nf = getenv("B2_NO_FUSED_ATTN");
    s->fused_attn = !(nf && nf[0] == '1');
    const char *npg = getenv("B2_NO_PREFILL_GEMM");
    s->use_pf_gemm = !(npg && npg[0] == '1');
    if (const char *gt = getenv("B2_GRAPH_TOKENS")) { s->multi_n = atoi(gt); if (s->multi_n < 1 || s->multi_n > 64) s->multi_n = 1; }
    if (m->tp > 1) {   // prompts of three tokens and more are evaluated in batches here too (prefill_batched: the row-parallel GEMMs end in an exchange)
        s->fused_attn = true;
        for (const LayerW &ly : m->layers)
            if (ly.split_qkv) { delete s; b2_set_error("tensor parallel sessions need q, k and v of a layer in one format"); return B2_EINVAL; }
    }
    int H = c.hidden, I = m->ffn(), qd = m->q_dim(), kvd = m->kv_dim(), hd = c.head_dim;
    uint64_t kv_elems = (uint64_t)c.n_layers * s->n_pages * m->kv_heads() * PAGE_TOKENS * hd;
    int rc = B2_OK;
    do {
        if ((rc = dev_alloc(s->allocs, (void **)&s->kcache, kv_elems * 2))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->vcache, kv_elems * 2))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->page_table, s->n_pages * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->state, sizeof(DevState)))) break;
        s->prompt_cap = max_ctx;
        s->out_cap = max_ctx;
        if ((rc = dev_alloc(s->allocs, (void **)&s->d_prompt, s->prompt_cap * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->d_out_tokens, s->out_cap * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->d_sampled, 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->d_ctx, (uint64_t)max_ctx * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->x, H * 4 * 2))) break;   // tensor parallel: the residual alternates between the halves
        if ((rc = dev_alloc(s->allocs, (void **)&s->qkv, (qd + 2 * kvd) * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->attn_out, qd * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->h, I * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->logits, (uint64_t)c.vocab * 4))) break;
        if ((rc = dev_alloc(s->allocs, &s->act_attn, a3_bytes(qd)))) break;
        if ((rc = dev_alloc(s->allocs, &s->act_h, a3_bytes(I)))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->pmax, 1024 * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->pidx, 1024 * 4))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->xsq, 1024 * 4))) break;
        if ((rc = dev_alloc(s->allocs, &s->attn_scratch, attn_scratch_bytes(m->heads(), hd)))) break;
        if ((rc = dev_alloc(s->allocs, &s->argmax_scratch, argmax_scratch_bytes()))) break;
        if ((rc = dev_alloc(s->allocs, &s->sample_scratch, sample_scratch_bytes(SAMPLE_MAX_K)))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->d_u, 16))) break;
        if ((rc = dev_alloc(s->allocs, (void **)&s->d_draw, 16))) break;
        if (s->use_pf_gemm) {
            const int64_t widest = std::max<int64_t>(std::max<int64_t>(H, I), qd);
            if ((rc = dev_alloc(s->allocs, (void **)&s->pf_x, (uint64_t)PF_MAX_TOKENS * H * 4))) break;
            if ((rc = dev_alloc(s->allocs, (void **)&s->pf_qkv, (uint64_t)PF_MAX_TOKENS * (qd + 2 * kvd) * 4))) break;
            if ((rc = dev_alloc(s->allocs, (void **)&s->pf_attn, (uint64_t)PF_MAX_TOKENS * qd * 4))) break;
            if ((rc = dev_alloc(s->allocs, (void **)&s->pf_h, (uint64_t)PF_MAX_TOKENS * I * 4))) break;
            if ((rc = dev_alloc(s->allocs, &s->pf_pb, pb_bytes(PF_MAX_TOKENS, widest)))) break;
            // split-K partial sums: up to 8 pieces of the H- or QKV-wide outputs, up to 3 of the gate/up pairs (prefill.cu: launch_prefill_gemm_qt)
            s->pf_part_floats = (uint64_t)PF_MAX_TOKENS * std::max<int64_t>(8 * std::max<int64_t>(H, qd + 2 * kvd), 3 * 2 * (int64_t)I);
            if ((rc = dev_alloc(s->allocs, (void **)&s->pf_part, s->pf_part_floats * 4))) break;
        }
    } while (0);
    if (rc) { b2_session_destroy(s); return rc; }
    // physical pages are handed out in reverse order so the page-table indirection is always exercised
    std::vector<int32_t> pt(s->n_pages);
    for (int p = 0; p < s->n_pages; ++p) pt[p] = s->n_pages - 1 - p;
    cudaError_t e = cudaMemcpy(s->page_table, pt.data(), pt.size() * 4, cudaMemcpyHostToDevice);
    // cache slots beyond the current position are read (and masked) by the tiled prompt attention: they must hold finite values
    if (e == cudaSuccess) e = cudaMemset(s->kcache, 0, kv_elems * 2);
    if (e == cudaSuccess) e = cudaMemset(s->vcache, 0, kv_elems * 2);
    if (e == cudaSuccess) e = cudaMemset(s->attn_scratch, 0, attn_scratch_bytes(m->heads(), hd));
    if (e == cudaSuccess) e = cudaMemset(s->argmax_scratch, 0, argmax_scratch_bytes());
    if (e == cudaSuccess) e = cudaMemset(s->sample_scratch, 0, sample_scratch_bytes(SAMPLE_MAX_K));
    if (e == cudaSuccess) e = cudaMemset(s->d_draw, 0, 16);
    if (e == cudaSuccess) e = cudaMemset(s->state, 0, sizeof(DevState));
    if (e == cudaSuccess) e = cudaMemset(s->act_attn, 0, a3_bytes(qd));   // padding blocks of the A3 vectors stay zero
    if (e == cudaSuccess) e = cudaMemset(s->act_h, 0, a3_bytes(I));
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev1);
    if (e != cudaSuccess) {
        b2_set_error("b2_session_create: %s", cudaGetErrorString(e));
        b2_session_destroy(s);
        return B2_ECUDA;
    }
    *out = s;
    return B2_OK;
}
extern "C" int b2_session_destroy(b2_session *s) {
    ApiScope _scope("b2_session_destroy");
    if (!s) return B2_OK;
    if (s->m && s->m->e) cudaSetDevice(s->m->e->device);
    drop_graphs(s);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->logits_log) cudaFree(s->logits_log);
    for (void *p : s->allocs) cudaFree(p);
    delete s;
    return B2_OK;
}
extern "C" int b2_session_reset(b2_session *s) {
    ApiScope _scope("b2_session_reset");
    B2_CHECK(s, B2_EINVAL, "NULL session");
    B2_CUDA(cudaSetDevice(s->m->e->device));
    B2_CUDA(cudaStreamSynchronize(s->m->e->stream));
    B2_CUDA(cudaMemset(s->state, 0, sizeof(DevState)));
    B2_CUDA(cudaMemset(s->d_draw, 0, 16));
    s->host_pos = 0;
    s->have_logits = false;
    uint64_t launches = s->stats.kernel_launches;
    s->stats = b2_stats{};
    s->stats.kernel_launches = launches;
    return B2_OK;
}
extern "C" int b2_session_pos(const b2_session *s, int *pos) {
    B2_CHECK(s && pos, B2_EINVAL, "NULL");
    *pos = s->host_pos;
    return B2_OK;
}
extern "C" int b2_session_stats(const b2_session *s, b2_stats *out) {
    B2_CHECK(s && out, B2_EINVAL, "NULL");
    *out = s->stats;
    return B2_OK;
}

static int ensure_log(b2_session *s, uint64_t floats) {
    if (floats <= s->logits_log_cap) return B2_OK;
    B2_CUDA(cudaStreamSynchronize(s->m->e->stream));
    if (s->logits_log) cudaFree(s->logits_log);
    s->logits_log = nullptr;
    s->logits_log_cap = 0;
    // graphs bake the log pointer: re-capture the ones that use it
    for (int g : {G_DECODE_LOG, G_FORCED, G_SAMPLE_LOG})
        for (int v = 0; v < ATTN_VARIANTS; ++v)
            if (s->graphs[g][v]) { cudaGraphExecDestroy(s->graphs[g][v]); s->graphs[g][v] = nullptr; }
    cudaError_t e = cudaMalloc((void **)&s->logits_log, floats * 4);
    if (e != cudaSuccess) { b2_set_error("logits log: cudaMalloc(%llu) failed: %s", (unsigned long long)floats * 4, cudaGetErrorString(e)); return B2_ENOMEM; }
    s->logits_log_cap = floats;
    return B2_OK;
}

static int set_step_counters(b2_session *s) {   // step = 0, prompt_idx = 0, keep pos
    DevState hs;
    hs.pos = s->host_pos;
    hs.step = 0;
    hs.cur_token = 0;
    hs.prompt_idx = 0;
    // cur_token is rewritten by the first kernel of every graph, so overwriting it here is harmless
    B2_CUDA(cudaMemcpyAsync(s->state, &hs, sizeof hs, cudaMemcpyHostToDevice, s->m->e->stream));
    return B2_OK;
}

// tensor-parallel sessions: a rank that waited in vain for a peer left a mark instead of hanging
static int tp_check(b2_session *s) {
    if (s->m->tp == 1) return B2_OK;
    unsigned int err = 0;
    B2_CUDA(cudaMemcpy(&err, s->m->e->tp.error(), 4, cudaMemcpyDeviceToHost));
    B2_CHECK(err == 0, B2_ENCCL, "tensor parallel exchange: rank %u did not arrive within 2 s (every rank must make the same calls with the same arguments)", err - 1);
    return B2_OK;
}

static int elapsed_ns(b2_session *s, uint64_t *ns) {
    float ms = 0.f;
    B2_CUDA(cudaEventSynchronize(s->ev1));
    B2_CUDA(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    *ns = (uint64_t)((double)ms * 1e6);
    return B2_OK;
}

// ------------------------------------------------------------------------------------ batched prompt evaluation
// n prompt tokens (already in s->d_prompt) at positions host_pos.. : every weight matrix is read once per batch of up to 128 tokens
// (tcgen05 GEMM, prefill.cu) instead of once per token; the LM head runs for the last token only, through the decode GEMV, and
// leaves logits and arg-max partials exactly where b2_decode expects them.
static int prefill_batched(b2_session *s, int n) {
    b2_model *m = s->m;
    const b2_arch_cfg &c = m->cfg;
    cudaStream_t st = m->e->stream;
    const int H = c.hidden, I = m->ffn(), qd = m->q_dim(), kvd = m->kv_dim(), hd = c.head_dim, sms = m->e->sm_count;
    const int64_t ldq = qd + 2 * kvd;
    const float escale = c.embed_scale ? sqrtf((float)H) : 1.0f;
    const uint64_t layer_kv = (uint64_t)s->n_pages * m->kv_heads() * PAGE_TOKENS * hd;
    int launches = 0;
    // tensor parallel: QKV and gate/up are column-parallel (this rank's rows, complete sums), O and down row-parallel: their split-K
    // partials are exchanged as tagged pairs (buffer 0: O rounds, buffer 1: down rounds, as in the decode pass) and the residual is added
    // to the sum over the ranks in rank order, so every rank holds the same batch of residuals
    const TpLink &L = m->e->tp;
    const int tpn = m->tp, rounds = 2 * c.n_layers;
    auto row_parallel = [&](const Mat &mt, int64_t kcols, int buf, int round, int mb, int T) -> int {
        int ks = 1;
        B2_CUDA(launch_prefill_gemm_qt(qt_of(mt.type), PF_OUT_BIAS, mt.packed, H, kcols, s->pf_pb, mb, T, nullptr, H, nullptr, 0, sms, s->pf_part, s->pf_part_floats, st,
                                       true, &ks));
        void *dst[TP_MAX_RANKS] = {};
        for (int p = 0; p < tpn; ++p) dst[p] = L.batch_slot(p, buf, L.rank);
        launch_pf_tp_exchange(s->pf_part, ks, T, mb, H, dst, L.batch_slot(L.rank, buf, 0), TpLink::batch_slot_pairs(), tpn, s->pf_x, s->pf_x, H, L.forwards(), rounds, round,
                              L.error(), st);
        return B2_OK;
    };
    // tokens per pass over the weights: 128 (B2_PF_TOKENS=64 / 32 / 16: experiments); a tensor-parallel group exchanges batches of TP_PF_TOKENS
    static int pf_tokens = 0;
    if (!pf_tokens) { const char *e = getenv("B2_PF_TOKENS"); const int v = e ? atoi(e) : 0; pf_tokens = (v == 16 || v == 32 || v == 64) ? v : PF_MAX_TOKENS; }
    const int step = tpn > 1 ? std::min(pf_tokens, TP_PF_TOKENS) : pf_tokens;
    for (int m0 = 0; m0 < n; m0 += step) {
        const int mb = n - m0 < step ? n - m0 : step, T = pf_batch(mb), pos0 = s->host_pos + m0;
        launch_prefill_embed(qt_of(m->embd.type), m->embd.packed, m->embd.cols, s->d_prompt + m0, mb, escale, s->pf_x, st); launches++;
        for (int l = 0; l < c.n_layers; ++l) {
            LayerW &ly = m->layers[l];
            B2_CUDA(launch_prefill_quant(PF_IN_NORM, s->pf_x, H, mb, T, H, ly.attn_norm, c.rms_eps, c.norm_plus_one, s->pf_pb, st));
            if (ly.split_qkv) {   // q, k, v in different formats (llama.cpp's Q4_K_M keeps attn_v in Q6_K): three GEMMs into the columns of the QKV batch
                int64_t row0 = 0;
                for (int part = 0; part < 3; ++part) {
                    const Mat &pm = ly.qkv_part[part];
                    B2_CUDA(launch_prefill_gemm_qt(qt_of(pm.type), PF_OUT_BIAS, pm.packed, pm.rows, H, s->pf_pb, mb, T, s->pf_qkv + row0, ldq,
                                                   ly.qkv_bias ? ly.qkv_bias + row0 : nullptr, 0, sms, s->pf_part, s->pf_part_floats, st));
                    row0 += pm.rows;
                    launches += part ? 1 : 0;
                }
            } else
            B2_CUDA(launch_prefill_gemm_qt(qt_of(ly.qkv.type), PF_OUT_BIAS, ly.qkv.packed, ly.qkv.rows, H, s->pf_pb, mb, T, s->pf_qkv, ldq, ly.qkv_bias, 0, sms, s->pf_part, s->pf_part_floats, st));
            B2_CUDA(launch_prefill_attention(s->pf_qkv, ldq, mb, m->heads(), m->kv_heads(), hd, pos0, m->rope_cos, m->rope_sin,
                                             s->kcache + l * layer_kv, s->vcache + l * layer_kv, s->page_table, s->pf_attn, qd, st));
            B2_CUDA(launch_prefill_quant(PF_IN_F32, s->pf_attn, qd, mb, T, qd, nullptr, 0.f, 0, s->pf_pb, st));
            if (tpn > 1) B2_TRY(row_parallel(ly.o, qd, 0, 2 * l, mb, T));
            else
            B2_CUDA(launch_prefill_gemm_qt(qt_of(ly.o.type), PF_OUT_RES, ly.o.packed, H, qd, s->pf_pb, mb, T, s->pf_x, H, s->pf_x, 0, sms, s->pf_part, s->pf_part_floats, st));
            B2_CUDA(launch_prefill_quant(PF_IN_NORM, s->pf_x, H, mb, T, H, ly.ffn_norm, c.rms_eps, c.norm_plus_one, s->pf_pb, st));
            B2_CUDA(launch_prefill_gemm_qt(qt_of(ly.gateup.type), PF_OUT_ACTMUL, ly.gateup.packed, 2 * (int64_t)I, H, s->pf_pb, mb, T, s->pf_h, I, nullptr, c.act, sms, s->pf_part, s->pf_part_floats, st));
            B2_CUDA(launch_prefill_quant(PF_IN_F32, s->pf_h, I, mb, T, I, nullptr, 0.f, 0, s->pf_pb, st));
            if (tpn > 1) B2_TRY(row_parallel(ly.down, I, 1, 2 * l + 1, mb, T));
            else
            B2_CUDA(launch_prefill_gemm_qt(qt_of(ly.down.type), PF_OUT_RES, ly.down.packed, H, I, s->pf_pb, mb, T, s->pf_x, H, s->pf_x, 0, sms, s->pf_part, s->pf_part_floats, st));
            launches += tpn > 1 ? 14 : 10;
        }
        if (tpn > 1) {   // the batch counts as one forward pass: the next pass's tags differ (every round of this one was seen complete by its gather)
            launch_tp_close(L.slot(L.rank, 1, 0), TP_MAX_VEC, tpn, H, L.forwards(), rounds, rounds - 1, false, L.error(), st);
            launches++;
        }
        if (m0 + mb == n) {   // LM head for the last token of the prompt
            B2_CUDA(cudaMemcpyAsync(s->x, s->pf_x + (int64_t)(mb - 1) * H, (size_t)H * 4, cudaMemcpyDeviceToDevice, st));
            launch_sumsq(s->x, H, s->xsq, st);
            launch_fill_partials(s->pmax, s->pidx, 1024, st);
            int grid = 0;
            B2_CUDA(launch_gemv_qt(qt_of(m->head.type), GEMV_IN_NORM, GEMV_OUT_LOGITS, m->head.packed, c.vocab, H, s->x, m->out_norm, c.rms_eps, c.norm_plus_one, nullptr,
                                s->logits, nullptr, nullptr, 0, s->pmax, s->pidx, &grid, nullptr, 0, s->xsq, 1, nullptr, st, false));
            s->head_grid = grid;
            launches += 3;
        }
    }
    B2_CUDA(cudaGetLastError());
    s->stats.kernel_launches += (uint64_t)launches;
    return B2_OK;
}

extern "C" int b2_prefill(b2_session *s, const int32_t *tokens, int n, float *logits_out) {
    ApiScope _scope("b2_prefill");
    B2_CHECK(s && tokens && n > 0, B2_EINVAL, "b2_prefill: bad arguments");
    B2_CHECK(s->host_pos + n <= s->max_ctx, B2_ECTX, "b2_prefill: %d + %d tokens exceed the context window %d", s->host_pos, n, s->max_ctx);
    int V = s->m->cfg.vocab;
    for (int i = 0; i < n; ++i) B2_CHECK(tokens[i] >= 0 && tokens[i] < V, B2_EINVAL, "b2_prefill: token %d out of range", tokens[i]);
    cudaStream_t st = s->m->e->stream;
    B2_CUDA(cudaSetDevice(s->m->e->device));
    B2_CUDA(cudaMemcpyAsync(s->d_prompt, tokens, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    B2_CUDA(cudaMemcpyAsync(s->d_ctx + s->host_pos, s->d_prompt, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    B2_TRY(set_step_counters(s));
    B2_CUDA(cudaEventRecord(s->ev0, st));
    if (s->use_pf_gemm && n >= 3) {
        B2_TRY(prefill_batched(s, n));
    } else {
        if (n > 1) B2_TRY(run_graph(s, G_PREFILL_BODY, n - 1, s->host_pos));
        B2_TRY(run_graph(s, G_PREFILL_LAST, 1, s->host_pos + n - 1));
    }
    B2_CUDA(cudaEventRecord(s->ev1, st));
    if (logits_out) B2_CUDA(cudaMemcpyAsync(logits_out, s->logits, (size_t)V * 4, cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    B2_TRY(tp_check(s));
    uint64_t ns = 0;
    B2_TRY(elapsed_ns(s, &ns));
    s->stats.prompt_eval_count += n;
    s->stats.prompt_eval_ns += ns;
    s->host_pos += n;
    s->have_logits = true;
    return B2_OK;
}

extern "C" int b2_decode(b2_session *s, int n_steps, const b2_sampler_cfg *sampler, int32_t *out_tokens, float *logits_out) {
    ApiScope _scope("b2_decode");
    B2_CHECK(s && out_tokens && n_steps > 0, B2_EINVAL, "b2_decode: bad arguments");
    B2_CHECK(s->have_logits, B2_EINVAL, "b2_decode: call b2_prefill first");
    const bool sampling = sampler && sampler->top_k > 1;
    if (sampling) {
        B2_CHECK(sampler->top_k <= SAMPLE_MAX_K, B2_EINVAL, "b2_decode: top_k %d > %d", sampler->top_k, SAMPLE_MAX_K);
        B2_CHECK(sampler->temperature > 0.f, B2_EINVAL, "b2_decode: temperature must be > 0 (use top_k <= 1 for greedy)");
        B2_CHECK(sampler->top_p > 0.f && sampler->top_p <= 1.0f, B2_EINVAL, "b2_decode: top_p must be in (0, 1]");
        B2_CHECK(sampler->repeat_penalty >= 0.f && sampler->repeat_last_n >= 0, B2_EINVAL, "b2_decode: repeat_penalty / repeat_last_n must not be negative");
    }
    B2_CHECK(s->host_pos + n_steps <= s->max_ctx, B2_ECTX, "b2_decode: %d + %d tokens exceed the context window %d", s->host_pos, n_steps, s->max_ctx);
    B2_CHECK(n_steps <= s->out_cap, B2_EINVAL, "b2_decode: n_steps %d > %d", n_steps, s->out_cap);
    cudaStream_t st = s->m->e->stream;
    int V = s->m->cfg.vocab;
    B2_CUDA(cudaSetDevice(s->m->e->device));
    if (logits_out) B2_TRY(ensure_log(s, (uint64_t)n_steps * V));
    B2_TRY(set_step_counters(s));
    B2_CUDA(cudaEventRecord(s->ev0, st));
    if (sampling) {
        // k / T / p are kernel arguments inside the captured graphs: new values -> new graphs.  The seed travels through device
        // memory; the draw counter continues across calls (reset by b2_session_reset).
        if (sampler->top_k != s->graph_sampler.top_k || sampler->temperature != s->graph_sampler.temperature ||
            sampler->top_p != s->graph_sampler.top_p || sampler->repeat_penalty != s->graph_sampler.repeat_penalty ||
            sampler->repeat_last_n != s->graph_sampler.repeat_last_n) {
            B2_CUDA(cudaStreamSynchronize(st));
            for (int g : {G_SAMPLE, G_SAMPLE_LOG})
                for (int v = 0; v < ATTN_VARIANTS; ++v)
                    if (s->graphs[g][v]) { cudaGraphExecDestroy(s->graphs[g][v]); s->graphs[g][v] = nullptr; }
            s->graph_sampler = *sampler;
        }
        B2_CUDA(cudaMemcpyAsync(s->d_draw, &sampler->seed, 8, cudaMemcpyHostToDevice, st));
        B2_TRY(run_graph(s, logits_out ? G_SAMPLE_LOG : G_SAMPLE, n_steps, s->host_pos));
    } else if (!logits_out && n_steps > 1 && tp_vocab_parallel(s->m)) {
        // all steps but the last with the LM head split over the ranks; the last one replicated, so that the call leaves complete
        // logits behind whatever the next call wants (sampling, logits_out)
        B2_TRY(run_graph(s, G_DECODE_VP, n_steps - 1, s->host_pos));
        B2_TRY(run_graph(s, G_DECODE, 1, s->host_pos + n_steps - 1));
    } else B2_TRY(run_graph(s, logits_out ? G_DECODE_LOG : G_DECODE, n_steps, s->host_pos));
    B2_CUDA(cudaEventRecord(s->ev1, st));
    B2_CUDA(cudaMemcpyAsync(out_tokens, s->d_out_tokens, (size_t)n_steps * 4, cudaMemcpyDeviceToHost, st));
    if (logits_out) B2_CUDA(cudaMemcpyAsync(logits_out, s->logits_log, (size_t)n_steps * V * 4, cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    B2_TRY(tp_check(s));
    uint64_t ns = 0;
    B2_TRY(elapsed_ns(s, &ns));
    s->stats.eval_count += n_steps;
    s->stats.eval_ns += ns;
    s->host_pos += n_steps;
    return B2_OK;
}

extern "C" int b2_forward_tokens(b2_session *s, const int32_t *tokens, int n, float *logits_out) {
    ApiScope _scope("b2_forward_tokens");
    B2_CHECK(s && tokens && n > 0, B2_EINVAL, "b2_forward_tokens: bad arguments");
    B2_CHECK(s->host_pos + n <= s->max_ctx, B2_ECTX, "b2_forward_tokens: context window %d exceeded", s->max_ctx);
    B2_CHECK(n <= s->prompt_cap, B2_EINVAL, "b2_forward_tokens: n too large");
    int V = s->m->cfg.vocab;
    for (int i = 0; i < n; ++i) B2_CHECK(tokens[i] >= 0 && tokens[i] < V, B2_EINVAL, "b2_forward_tokens: token %d out of range", tokens[i]);
    cudaStream_t st = s->m->e->stream;
    B2_CUDA(cudaSetDevice(s->m->e->device));
    B2_TRY(ensure_log(s, (uint64_t)n * V));
    B2_CUDA(cudaMemcpyAsync(s->d_prompt, tokens, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    B2_CUDA(cudaMemcpyAsync(s->d_ctx + s->host_pos, s->d_prompt, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    B2_TRY(set_step_counters(s));
    B2_TRY(run_graph(s, G_FORCED, n, s->host_pos));
    if (logits_out) B2_CUDA(cudaMemcpyAsync(logits_out, s->logits_log, (size_t)n * V * 4, cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    B2_TRY(tp_check(s));
    s->host_pos += n;
    s->have_logits = true;
    return B2_OK;
}

// ------------------------------------------------------------------------------------ host random init
extern "C" int b2_randinit_tensor(uint64_t seed, uint32_t tensor_id, int64_t rows, int64_t cols, int type, float sigma, float offset,
                                  int64_t row0, int64_t nrows, void *out, uint64_t out_bytes) {
    B2_CHECK(out && rows > 0 && row0 >= 0 && nrows >= 0 && row0 + nrows <= rows, B2_EINVAL, "b2_randinit_tensor: bad arguments");
    uint64_t key = ri_tensor_key(seed, tensor_id);
    if (cols == 0 || type == B2_T_F32) {
        int64_t c = cols == 0 ? 1 : cols;
        int64_t n = nrows * c;
        B2_CHECK(out_bytes == (uint64_t)n * 4, B2_EINVAL, "b2_randinit_tensor: out_bytes %llu, expected %llu", (unsigned long long)out_bytes, (unsigned long long)n * 4);
        float *o = (float *)out;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) o[i] = ri_add(ri_mul(ri_gauss(key, (uint64_t)(row0 * c + i)), sigma), offset);
        return B2_OK;
    }
    B2_CHECK(type == B2_T_Q4_0, B2_EINVAL, "b2_randinit_tensor: type %d not supported", type);
    B2_CHECK(cols % QK == 0, B2_EINVAL, "b2_randinit_tensor: cols %% 32");
    int64_t nb = cols / QK, n = nrows * nb;
    B2_CHECK(out_bytes == (uint64_t)n * 18, B2_EINVAL, "b2_randinit_tensor: out_bytes %llu, expected %llu", (unsigned long long)out_bytes, (unsigned long long)n * 18);
    uint8_t *o = (uint8_t *)out;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        float x[QK];
        ri_gauss_block(key, (uint64_t)(row0 * nb + i) * QK, sigma, x);
        uint16_t dbits;
        uint32_t w[4];
        ri_quant_q4_0(x, dbits, w);
        uint8_t *p = o + i * 18;
        p[0] = (uint8_t)(dbits & 0xFF);
        p[1] = (uint8_t)(dbits >> 8);
        memcpy(p + 2, w, 16);
    }
    return B2_OK;
}

// ------------------------------------------------------------------------------------ device helpers
extern "C" int b2_dev_alloc(uint64_t nbytes, void **d_out) {
    ApiScope _scope("b2_dev_alloc");
    B2_CHECK(d_out, B2_EINVAL, "NULL");
    B2_CUDA(cudaMalloc(d_out, nbytes ? nbytes : 16));
    return B2_OK;
}
extern "C" int b2_dev_free(void *p) {
    ApiScope _scope("b2_dev_free");
    if (p) B2_CUDA(cudaFree(p));
    return B2_OK;
}
extern "C" int b2_dev_upload(void *d_dst, const void *src, uint64_t n) {
    ApiScope _scope("b2_dev_upload");
    B2_CUDA(cudaMemcpy(d_dst, src, n, cudaMemcpyHostToDevice));
    return B2_OK;
}
extern "C" int b2_dev_download(void *dst, const void *d_src, uint64_t n) {
    ApiScope _scope("b2_dev_download");
    B2_CUDA(cudaMemcpy(dst, d_src, n, cudaMemcpyDeviceToHost));
    return B2_OK;
}
extern "C" int b2_dev_sync(void) {
    ApiScope _scope("b2_dev_sync");
    B2_CUDA(cudaDeviceSynchronize());
    return B2_OK;
}
