// q4k_common.cuh — block layouts and per-thread arithmetic shared by every kernel of the path.
//
// Layouts are the reference's bytes, unchanged ("L/" = external/llama.cpp/ in the reference):
//   block_q4_K  L/ggml/src/ggml-common.h:291-306   144 B = one 16 B header chunk + eight 16 B nibble chunks
//   block_q8_K  L/ggml/src/ggml-common.h:338-344   292 B = f32 d, 256 x int8, 16 x int16 bsums
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace q4k {

constexpr int QK_K      = 256;
constexpr int BLK_Q4K   = 144;
constexpr int BLK_Q8K   = 292;
constexpr int Q8K_QS    = 4;     // offsetof(block_q8_K, qs)
constexpr int Q8K_BSUMS = 260;   // offsetof(block_q8_K, bsums)

struct __align__(16) block_q4_K {
    __half  d;
    __half  dmin;
    uint8_t scales[12];
    uint8_t qs[128];
};
static_assert(sizeof(block_q4_K) == BLK_Q4K, "block_q4_K must stay 144 bytes");

// streaming 16-byte load of weights: read once, keep out of L1
__device__ __forceinline__ uint4 ldg_stream16(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// The 16 B header chunk of a block_q4_K: word0 = {d, dmin} as two fp16, words 1..3 = 12 packed bytes.
// Unpacks the 6-bit scales and mins into four words of four bytes each:
//   sc[0..3] = u.x, sc[4..7] = u.y, m[0..3] = u.z, m[4..7] = u.w
// (same result as get_scale_min_k4, L/ggml/src/ggml-quants.c:703-710, for j = 0..7)
__device__ __forceinline__ uint4 unpack_scales_mins(const uint4 hdr) {
    const uint32_t s0 = hdr.y, s1 = hdr.z, s2 = hdr.w;
    uint4 u;
    u.x = s0 & 0x3f3f3f3fu;
    u.y = (s2 & 0x0f0f0f0fu) | (((s0 >> 6) & 0x03030303u) << 4);
    u.z = s1 & 0x3f3f3f3fu;
    u.w = ((s2 >> 4) & 0x0f0f0f0fu) | (((s1 >> 6) & 0x03030303u) << 4);
    return u;
}

// bytes (2*pair, 2*pair+1) of the 8-byte sequence {lo_word, hi_word}, as one 16-bit value (pair = 0..3)
__device__ __forceinline__ uint32_t byte_pair(uint32_t lo_word, uint32_t hi_word, int pair) {
    const uint32_t w = (pair & 2) ? hi_word : lo_word;
    return (w >> ((pair & 1) * 16)) & 0xffffu;
}

__device__ __forceinline__ float header_d(const uint4 hdr) {
    return __half2float(__ushort_as_half((unsigned short)(hdr.x & 0xffffu)));
}
__device__ __forceinline__ float header_dmin(const uint4 hdr) {
    return __half2float(__ushort_as_half((unsigned short)(hdr.x >> 16)));
}

// nearest_int of the reference (L/ggml/src/ggml-quants.c:444-449): the rounding IS the float add.
__device__ __forceinline__ int nearest_int(float f) {
    const float shifted = __fadd_rn(f, 12582912.0f);
    return (int)(__float_as_uint(shifted) & 0x7fffffu) - 0x400000;
}

}  // namespace q4k
