// randinit.cuh — deterministic random initialiser shared by host (b2_randinit_tensor) and device
// (b2_model_create_random).  Integer-exact recipe, restated independently in oracle/randinit.py:
//   key = splitmix64(seed * 0x100000001B3 + tensor_id);  h = splitmix64(key + i)
//   S = sum of the four 16-bit fields of h;  g = float(S - 131070) * GSCALE;  value = g * sigma (+ offset)
// followed by the ggml Q4_0 quantiser (quantize_row_q4_0_ref) with every float op individually rounded,
// so host and device produce identical bytes.  The host side must be compiled with -ffp-contract=off.
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>

#ifdef __CUDACC__
#define RI_HD __host__ __device__ __forceinline__
#else
#define RI_HD inline
#endif

#define RI_GSCALE 2.6429157e-05f

RI_HD float ri_mul(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fmul_rn(a, b);
#else
    volatile float r = a * b;
    return r;
#endif
}
RI_HD float ri_add(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fadd_rn(a, b);
#else
    volatile float r = a + b;
    return r;
#endif
}
RI_HD float ri_div(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fdiv_rn(a, b);
#else
    volatile float r = a / b;
    return r;
#endif
}

RI_HD uint64_t ri_splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
RI_HD uint64_t ri_tensor_key(uint64_t seed, uint32_t tensor_id) { return ri_splitmix64(seed * 0x100000001B3ull + tensor_id); }

RI_HD float ri_gauss(uint64_t key, uint64_t i) {
    uint64_t h = ri_splitmix64(key + i);
    int s = (int)(h & 0xFFFF) + (int)((h >> 16) & 0xFFFF) + (int)((h >> 32) & 0xFFFF) + (int)(h >> 48);
    return ri_mul((float)(s - 131070), RI_GSCALE);
}
RI_HD void ri_gauss_block(uint64_t key, uint64_t i0, float sigma, float *x /*32*/) {
    for (int j = 0; j < 32; ++j) x[j] = ri_mul(ri_gauss(key, i0 + j), sigma);
}

// float -> IEEE half bits, round-to-nearest-even (host fallback; device uses the intrinsic)
RI_HD uint16_t ri_f32_to_f16_bits(float f) {
#ifdef __CUDA_ARCH__
    return __half_as_ushort(__float2half_rn(f));
#else
    uint32_t x;
    memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t absx = x & 0x7FFFFFFFu;
    if (absx >= 0x7F800000u) return (uint16_t)(sign | (absx > 0x7F800000u ? 0x7E00u : 0x7C00u));
    if (absx >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);  // rounds to inf
    if (absx < 0x33000001u) return (uint16_t)sign;               // rounds to zero (< 2^-25, or == 2^-25 tie to even 0)
    int e = (int)(absx >> 23) - 127;
    uint32_t m = (absx & 0x7FFFFFu) | 0x800000u;
    int shift;
    uint32_t base;
    if (e < -14) { shift = 13 + (-14 - e); base = 0; }           // subnormal half
    else { shift = 13; base = (uint32_t)(e + 15) << 10; m &= 0x7FFFFFu; }
    uint32_t q = m >> shift, rem = m & ((1u << shift) - 1), halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (q & 1))) q++;
    return (uint16_t)(sign | (base + q));
#endif
}

// ggml quantize_row_q4_0_ref on one block of 32: returns f16 bits of d and 4 words of nibbles
RI_HD void ri_quant_q4_0(const float *x, uint16_t &dbits, uint32_t *w /*4*/) {
    float amax = 0.f, mx = 0.f;
    for (int j = 0; j < 32; ++j) {
        float v = x[j], a = fabsf(v);
        if (amax < a) { amax = a; mx = v; }
    }
    float d = ri_div(mx, -8.0f);
    float id = d != 0.f ? ri_div(1.0f, d) : 0.f;
    dbits = ri_f32_to_f16_bits(d);
    for (int k = 0; k < 4; ++k) w[k] = 0;
    for (int j = 0; j < 16; ++j) {
        int q0 = (int)(int8_t)ri_add(ri_mul(x[j], id), 8.5f);
        int q1 = (int)(int8_t)ri_add(ri_mul(x[16 + j], id), 8.5f);
        if (q0 > 15) q0 = 15;
        if (q1 > 15) q1 = 15;
        uint32_t byte = (uint32_t)(q0 & 0xF) | ((uint32_t)(q1 & 0xF) << 4);
        w[j >> 2] |= byte << (8 * (j & 3));
    }
}
