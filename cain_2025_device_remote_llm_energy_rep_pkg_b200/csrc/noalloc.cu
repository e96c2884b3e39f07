// noalloc.cu — greedy sampler entry point for callers that must not allocate device memory.
//
// b2_op_argmax needs a little scratch (per-block partial maxima and a ticket counter) and gets it from cudaMalloc on
// every call.  Some users of this library cannot afford that: the server thread that samples a token while a measured
// request is running (cudaMalloc/cudaFree synchronise the whole device and would stall the decode whose time and energy
// the experiment records), and processes that sized the KV cache and weights to fill the card, where a late
// cudaMalloc can simply fail.  For them the same kernel runs here on scratch that lives in the library image itself.
#include "kernels.cuh"
#include <mutex>

#define NOALLOC_MAX_DEVICES 64
#define NOALLOC_SCRATCH_BYTES 2048   // >= argmax_scratch_bytes() (128 partial values + 128 indices + the counter)

// Lives in every GPU context that loads this library, from load until exit.  In host code this name is only a
// handle for cudaGetSymbolAddress, never a pointer a kernel could use.
static __device__ __align__(256) unsigned char g_argmax_scratch[NOALLOC_SCRATCH_BYTES];

static std::mutex g_mu;
static void *g_scratch_ptr[NOALLOC_MAX_DEVICES];   // device address of g_argmax_scratch, per device ordinal

extern "C" int b2_op_argmax_noalloc(const float *d_logits, int64_t n, int32_t *d_out_index) {
    ApiScope _scope("b2_op_argmax_noalloc");
    B2_CHECK(d_logits && d_out_index, B2_EINVAL, "b2_op_argmax_noalloc: NULL pointer");
    B2_CHECK(n > 0, B2_EINVAL, "b2_op_argmax_noalloc: n must be positive");
    B2_CHECK(n <= B2_ARGMAX_NOALLOC_MAX_N, B2_EINVAL, "b2_op_argmax_noalloc: n = %lld exceeds the supported maximum of %d values; use b2_op_argmax",
             (long long)n, B2_ARGMAX_NOALLOC_MAX_N);
    B2_CHECK(argmax_scratch_bytes() <= NOALLOC_SCRATCH_BYTES, B2_EINVAL, "b2_op_argmax_noalloc: built-in scratch too small");
    int dev = -1;
    B2_CUDA(cudaGetDevice(&dev));
    B2_CHECK(dev >= 0 && dev < NOALLOC_MAX_DEVICES, B2_EINVAL, "b2_op_argmax_noalloc: device ordinal %d not supported", dev);
    void *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_scratch_ptr[dev]) {   // first call on this device (it is the current one): resolve the symbol once
            void *p = nullptr;
            B2_CUDA(cudaGetSymbolAddress(&p, g_argmax_scratch));
            g_scratch_ptr[dev] = p;
        }
        scratch = g_scratch_ptr[dev];
    }
    // the array keeps whatever the previous call left: the ticket counter has to start at zero
    B2_CUDA(cudaMemsetAsync((char *)scratch + ARGMAX_BLOCKS * 8, 0, 4, 0));
    launch_argmax(d_logits, n, d_out_index, scratch, 0);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
