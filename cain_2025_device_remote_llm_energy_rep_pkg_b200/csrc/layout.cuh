// layout.cuh — HBM data layouts of the decode path.
//
// Activation vector ("A3"), n elements, nbp = blocks of 32 padded up to a multiple of 4 (nbp/4 block-groups):
//   [0, nbp*96)        int8 digits: block-group bg at bg*384, digit plane c (0..2) at +c*128, inside a plane
//                      [t 0..3][j 0..3] 8 bytes = elements 4t..4t+3 then 16+4t..16+4t+3 of block 4*bg+j, so lane
//                      (g = c, t) of an MMA warp reads the B fragments of four consecutive blocks as 32 contiguous bytes
//   [nbp*96, nbp*100)  float scale[nbp] = 2^(e-19)  (0 for padding blocks)
//   X = d2*16384 + d1*128 + d0 (balanced base-128 digits), x_hat = X * scale.
//
// Q4_0 weight matrix [rows, cols], packed for streaming: rows in tiles of 16, blocks padded to nbp (x4):
//   tile T at T*nbp*288.  Inside a tile the K dimension is cut into chunks of KC=64 blocks (the last may be
//   shorter); chunk ck starts at ck*KC*288 and holds cb blocks as
//     nibbles  cb*256 B : block-group j (4 blocks) at j*1024: [half hf][g 0..7][t 0..3] uint4 =
//                         word t of blocks 4j..4j+3 of tile row hf*8+g; nibbles stored as two's complement
//                         of (q-8) (GGML nibble XOR 8), GGML nibble order kept (low nibble = element i,
//                         high nibble = element i+16 of the block)
//     scales   cb*32 B  : [block-group][g][j] half2 = (d of row g, d of row g+8) of block 4*bg+j
//   One chunk is one contiguous cp.async.bulk; one warp-wide LDS.128 of the nibble area feeds the A fragments
//   of four mma.m16n8k32 instructions.
#pragma once
#include "common.cuh"

#define TILE_ROWS 16
#define KC 64                 // blocks per chunk
#define BLOCK_BYTES 288       // 16 rows * 16 B nibbles + 16 rows * 2 B scales, per 32-wide block column of a tile
#define CHUNK_BYTES (KC * BLOCK_BYTES)

__host__ __device__ inline int64_t pad4(int64_t nb) { return (nb + 3) & ~(int64_t)3; }
__host__ __device__ inline int64_t nblocks_padded(int64_t cols) { return pad4((cols + QK - 1) / QK); }
__host__ __device__ inline int64_t ntiles(int64_t rows) { return (rows + TILE_ROWS - 1) / TILE_ROWS; }
__host__ __device__ inline uint64_t packed_q4_0_bytes(int64_t rows, int64_t cols) {
    return (uint64_t)ntiles(rows) * nblocks_padded(cols) * BLOCK_BYTES;
}
__host__ __device__ inline uint64_t a3_bytes(int64_t n) { return (uint64_t)nblocks_padded(n) * 100; }
// byte offset of (block blk, digit plane c, element e 0..31) inside the digit area
__host__ __device__ inline int64_t a3_off(int64_t blk, int c, int e) {
    return (blk >> 2) * 384 + c * 128 + ((((e & 15) >> 2) * 4 + (blk & 3)) << 3) + ((e >> 4) << 2) + (e & 3);
}

// location of (row, block) inside the packed matrix: returns byte offsets of the 4-byte nibble word t and of the scale
struct PackedLoc {
    uint64_t word0;   // byte offset of word t=0 for this (row, block); word t is at word0 + t*16
    uint64_t scale;   // byte offset of the half scale
};
__host__ __device__ inline PackedLoc packed_loc(int64_t row, int64_t blk, int64_t nbp) {
    int64_t T = row / TILE_ROWS, r = row % TILE_ROWS, hf = r >> 3, g = r & 7;
    int64_t ck = blk / KC, bl = blk % KC;
    int64_t cb = (nbp - ck * KC) < KC ? (nbp - ck * KC) : KC;
    uint64_t chunk = (uint64_t)T * nbp * BLOCK_BYTES + (uint64_t)ck * KC * BLOCK_BYTES;
    PackedLoc p;
    p.word0 = chunk + (uint64_t)(bl >> 2) * 1024 + (uint64_t)(hf * 8 + g) * 64 + (uint64_t)(bl & 3) * 4;
    p.scale = chunk + (uint64_t)cb * 256 + (uint64_t)(bl >> 2) * 128 + (uint64_t)g * 16 + (uint64_t)(bl & 3) * 4 + (uint64_t)hf * 2;
    return p;
}

// ---- Q4_K matrices (GGML super-blocks of 256: f16 d, f16 dmin, eight 6-bit (scale, min) pairs, 128 B of nibbles;
// w = d*sc*q - dmin*m).  Same tiles, chunks and nibble area as Q4_0 (nibbles stored as q-8, so the int8 MMA path is shared and
// w = A*(q-8) + (8A - B) with A = d*sc, B = dmin*m); the scale area of a block-group grows from 128 to 192 bytes:
//     [g 0..7] 16 B : word j = (sc of row g, m of row g, sc of row g+8, m of row g+8) of block 4*bg+j
//     [g 0..7]  8 B : half2 (d, dmin) of row g, half2 (d, dmin) of row g+8 -- of the super-block this block-group lies in
// 304 bytes per 32-wide block column of a tile = 4.75 bit per weight (GGML: 4.5).
#define BLOCK_BYTES_K 304
__host__ __device__ inline uint64_t packed_q4_k_bytes(int64_t rows, int64_t cols) {
    return (uint64_t)ntiles(rows) * nblocks_padded(cols) * BLOCK_BYTES_K;
}
struct PackedLocK {
    uint64_t word0;   // as PackedLoc
    uint64_t scm;     // byte offset of sc (m follows at +1)
    uint64_t dm;      // byte offset of the half2 (d, dmin)
};
__host__ __device__ inline PackedLocK packed_loc_k(int64_t row, int64_t blk, int64_t nbp) {
    int64_t T = row / TILE_ROWS, r = row % TILE_ROWS, hf = r >> 3, g = r & 7;
    int64_t ck = blk / KC, bl = blk % KC;
    int64_t cb = (nbp - ck * KC) < KC ? (nbp - ck * KC) : KC;
    uint64_t chunk = (uint64_t)T * nbp * BLOCK_BYTES_K + (uint64_t)ck * KC * BLOCK_BYTES_K;
    uint64_t sc = chunk + (uint64_t)cb * 256 + (uint64_t)(bl >> 2) * 192;
    PackedLocK p;
    p.word0 = chunk + (uint64_t)(bl >> 2) * 1024 + (uint64_t)(hf * 8 + g) * 64 + (uint64_t)(bl & 3) * 4;
    p.scm = sc + (uint64_t)g * 16 + (uint64_t)(bl & 3) * 4 + (uint64_t)hf * 2;
    p.dm = sc + 128 + (uint64_t)g * 8 + (uint64_t)hf * 4;
    return p;
}

// ---- Q6_K matrices (GGML super-blocks of 256: 4 low bits + 2 high bits per weight, sixteen int8 scales for 16-element
// sub-blocks, f16 d; w = d*sc*(q-32)).  llama.cpp keeps output.weight (and a tied token_embd.weight) in this format even in
// "Q4_0" files.  Same tiles and chunks; per chunk of cb blocks:
//     low nibbles  cb*256 B : as Q4_0, nibble = (q & 15) - 8 in two's complement
//     high bits    cb*128 B : block-group j at j*512: [lane = g*4+t] uint4 = {row g blocks 4j,4j+1 | row g blocks 4j+2,4j+3 |
//                             row g+8 blocks 4j,4j+1 | row g+8 blocks 4j+2,4j+3}; 16 bits per (row, block): 2 bits per element
//                             4t..4t+3 (low byte) and 16+4t..16+4t+3 (high byte) -- exactly the elements of this lane's A fragments
//     scales       per block-group 160 B: [g] 16 B: word j = int8 (sc_lo, sc_hi of row g, sc_lo, sc_hi of row g+8) of block 4*bg+j
//                             (sc_lo: elements 0..15, sc_hi: 16..31), then [g] 4 B: half2 (d of row g, d of row g+8)
// 424 bytes per 32-wide block column of a tile = 6.625 bit per weight (GGML: 6.5625).
#define BLOCK_BYTES_6 424
__host__ __device__ inline uint64_t packed_q6_k_bytes(int64_t rows, int64_t cols) {
    return (uint64_t)ntiles(rows) * nblocks_padded(cols) * BLOCK_BYTES_6;
}
struct PackedLoc6 {
    uint64_t word0;   // low-nibble word t = 0 (word t at +16*t), as PackedLoc
    uint64_t hi;      // byte offset of the 16-bit high-bit field of lane t = 0 for this (row, block); lane t at +16*t
    uint64_t sc;      // byte offset of sc_lo (sc_hi at +1)
    uint64_t d;       // byte offset of the half d
};
__host__ __device__ inline PackedLoc6 packed_loc_6(int64_t row, int64_t blk, int64_t nbp) {
    int64_t T = row / TILE_ROWS, r = row % TILE_ROWS, hf = r >> 3, g = r & 7;
    int64_t ck = blk / KC, bl = blk % KC;
    int64_t cb = (nbp - ck * KC) < KC ? (nbp - ck * KC) : KC;
    uint64_t chunk = (uint64_t)T * nbp * BLOCK_BYTES_6 + (uint64_t)ck * KC * BLOCK_BYTES_6;
    uint64_t sa = chunk + (uint64_t)cb * 384 + (uint64_t)(bl >> 2) * 160;
    PackedLoc6 p;
    p.word0 = chunk + (uint64_t)(bl >> 2) * 1024 + (uint64_t)(hf * 8 + g) * 64 + (uint64_t)(bl & 3) * 4;
    p.hi = chunk + (uint64_t)cb * 256 + (uint64_t)(bl >> 2) * 512 + (uint64_t)g * 64 + (uint64_t)hf * 8 + (uint64_t)(bl & 3) * 2;
    p.sc = sa + (uint64_t)g * 16 + (uint64_t)(bl & 3) * 4 + (uint64_t)hf * 2;
    p.d = sa + 128 + (uint64_t)g * 4 + (uint64_t)hf * 2;
    return p;
}

#ifdef __CUDACC__
// store one activation block (warp-cooperative, lane = element) into an A3 vector of n elements
__device__ __forceinline__ void a3_store_block(void *act, int64_t n, int64_t blk, int lane, int X, float scale) {
    int d0, d1, d2;
    act_digits(X, d0, d1, d2);
    uint8_t *dst = (uint8_t *)act + a3_off(blk, 0, lane);
    dst[0] = (uint8_t)d0;
    dst[128] = (uint8_t)d1;
    dst[256] = (uint8_t)d2;
    if (lane == 0) ((float *)((uint8_t *)act + nblocks_padded(n) * 96))[blk] = scale;
}
#endif
