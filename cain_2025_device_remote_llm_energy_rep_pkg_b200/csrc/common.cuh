// common.cuh — shared device helpers and host error plumbing for libb2llm.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include "../../include/b2llm.h"

// ---------------------------------------------------------------- host error plumbing
void b2_set_error(const char *fmt, ...);
#define B2_CUDA(expr)                                                                         \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            b2_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return B2_ECUDA;                                                                  \
        }                                                                                     \
    } while (0)
#define B2_CHECK(cond, code, ...)                                                             \
    do {                                                                                      \
        if (!(cond)) { b2_set_error(__VA_ARGS__); return (code); }                            \
    } while (0)
#define B2_TRY(expr)                                                                          \
    do { int _r = (expr); if (_r != B2_OK) return _r; } while (0)

// Every extern "C" entry point opens an ApiScope: it swallows (and, with B2_DEBUG_STALE=1, reports) a
// non-sticky CUDA error left behind by an earlier call so that it is not blamed on this one.
struct ApiScope {
    const char *name;
    explicit ApiScope(const char *n);
    ~ApiScope();
};

// ---------------------------------------------------------------- block geometry
#define QK        32      // weights per Q4_0 / Q8_0 block and per activation block
#define QK_K      256     // weights per K-quant super-block
#define ACT_BITS  20      // |X| <= 2^20 on the activation grid
#define PAGE_TOKENS 16    // tokens per KV page
#define B2_MAX_DEVICES 64 // GPUs one process may drive (per-device caches of function attributes)

#ifdef __CUDACC__
// ---------------------------------------------------------------- warp helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// dot of 4 unsigned bytes (weights 0..15) with 4 signed bytes (activation digits)
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp4a_ss(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {  // weights: read once, keep out of L1
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// Activation grid, one block of 32 per warp: lane holds x; returns X (integer on the grid) and the
// block's scale 2^(e-19), e = max(floor(log2 max|x|), -100).  Matches oracle/model.py:fxp_round.
__device__ __forceinline__ int act_grid_lane(float x, float &scale_out) {
    float m = warp_max(fabsf(x));
    int e = (int)((__float_as_uint(m) >> 23) & 0xFF) - 127;   // floor(log2 m) for normal m
    e = max(e, -100);
    // scale = 2^(e-19), inv = 2^(19-e): both exact powers of two, exponent fields in range
    scale_out = __uint_as_float((uint32_t)(e - (ACT_BITS - 1) + 127) << 23);
    float inv = __uint_as_float((uint32_t)((ACT_BITS - 1) - e + 127) << 23);
    return __float2int_rn(x * inv);
}
// split X into balanced base-128 digits
__device__ __forceinline__ void act_digits(int X, int &d0, int &d1, int &d2) {
    d0 = ((X + 64) & 127) - 64;
    int X1 = (X - d0) >> 7;
    d1 = ((X1 + 64) & 127) - 64;
    d2 = (X1 - d1) >> 7;
}
// arg-max with the lowest index on ties (greedy sampler)
__device__ __forceinline__ void arg_better(float &v, int &i, float ov, int oi) {
    if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}
__device__ __forceinline__ void warp_argmax(float &v, int &i) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, v, o);
        int oi = __shfl_xor_sync(0xffffffffu, i, o);
        arg_better(v, i, ov, oi);
    }
}
#endif
