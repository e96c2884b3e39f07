// mega.cuh — the persistent token kernel (mega.cu): plan layout and launcher.
#pragma once
#include "gemv_dev.cuh"
#include "attn_dev.cuh"

#define MEGA_MAX_LAYERS 40

// what differs from layer to layer: weights and the layer's slice of the KV cache.  Everything else about a stage is the same in
// every layer and sits at a fixed constant-bank address (an instruction operand, no register), like a kernel parameter.
struct MegaLayer {
    const uint8_t *qkv_w, *o_w, *gateup_w, *down_w;
    const float *attn_norm, *ffn_norm, *qkv_bias;
    __half *kc, *vc;
};
// Everything the kernel needs that does not change between launches of one session; lives in __constant__ memory while
// that session is the one being run (re-uploaded, stream-ordered, when another session takes over).
struct MegaPlan {
    int n_layers, n_embd, vocab, n_part;   // n_part: per-CTA partials (= the grid the plan was built for)
    const uint8_t *embd_w;                 // packed token-embedding matrix
    int embd_nbp;
    float embd_scale;
    float *x;                              // residual stream
    float *xsq;                            // partial sums of squares of x (first half; the o-projection writes the second)
    const float *logits;
    const float *pmax;
    const int *pidx;
    GemvArgs qkv, o, gateup, down, head;   // per-stage argument blocks (layer 0's pointers; the kernel patches in the layer's)
    AttnArgs attn;
    MegaLayer layers[MEGA_MAX_LAYERS];
};
enum { MEGA_PROMPT = 0, MEGA_DECODE = 1 };
struct MegaRun {
    int mode;            // MEGA_PROMPT: tokens come from `prompt`; MEGA_DECODE: arg-max of the previous LM head
    int n_tokens;
    int head_from;       // tokens with index >= head_from run the LM head
    int pos0;            // position of the first token
    const int *prompt;
    int *out_tokens;     // MEGA_DECODE: the sampled tokens
    float *logits_log;   // optional [n_tokens][vocab]: DECODE: the logits each token was sampled from; PROMPT: the logits each token produced
    DevState *state;     // pos/step are left as the per-token graphs would leave them
};

// grid the plan must be built for (one CTA per SM if the kernel fits co-resident, as it must for a cooperative launch)
cudaError_t mega_grid(size_t smem, int *grid_out);
// `owner` identifies the plan (session); the constant copy is refreshed when it differs from the one last launched on this device
cudaError_t launch_mega(const MegaPlan *host_plan, const void *owner, const MegaRun &run, int grid, size_t smem, cudaStream_t st);
// forget `owner` (session destroyed: a new session may reuse the address)
void mega_forget(const void *owner);
