// q4k_common.cuh — block layouts and per-thread arithmetic shared by every kernel of the path.
//
// Layouts are the reference's bytes, unchanged ("L/" = external/llama.cpp/ in the reference):
//   block_q4_K  L/ggml/src/ggml-common.h:291-306   144 B = one 16 B header chunk + eight 16 B nibble chunks
//   block_q8_K  L/ggml/src/ggml-common.h:338-344   292 B = f32 d, 256 x int8, 16 x int16 bsums
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#ifdef Q4K_TIMING          // tools/micro/mma_ticks.cu: per-phase clock64 stamps of CTA 0 (development only)
__device__ long long g_ticks[32];
#define TICK(i) do { if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) { long long t_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_)); g_ticks[i] = t_; } } while (0)
#define GTICK(i) do { if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) { long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_ticks[i] = t_; } } while (0)
// QT: stamp after the value `dep` exists (warp 0 of CTA 0)
#define QT(i, dep) do { if (blockIdx.x == 0 && threadIdx.x < 32) { asm volatile("" ::"r"(dep)); TICK(i); } } while (0)
#else
#define TICK(i) do {} while (0)
#define GTICK(i) do {} while (0)
#define QT(i, dep) do {} while (0)
#endif

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

// Coherent 16-byte load of activations.  The quantizer kernels run under programmatic dependent launch: their lifetime overlaps
// the kernel (or the peer GPU) that produces x, so the read-only contract of ld.global.nc — which nvcc picks for a
// `const __restrict__` pointer — does not hold, even though every load sits behind griddepcontrol.wait.
__device__ __forceinline__ float4 ld_act16(const float *p) {
    float4 v;
    asm volatile("ld.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

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

// One 256-float block of quantize_row_q8_K_ref (L/ggml/src/ggml-quants.c:2555-2592) on one warp.  Lane l
// holds floats [4l, 4l+4) and [128+4l, 128+4l+4) in v[0..3], v[4..7].  Every float operation is an explicit
// single-rounding intrinsic (no FMA contraction possible):
//   absmax : max of |x| as unsigned bit patterns (non-negative floats order like integers): per lane, then a
//            five-step butterfly of single shuffles.  "First occurrence wins" of the scalar loop: elements of the
//            first half (index 4l + i) precede all of the second half (128 + 4l + i) and are ordered by lane, so the
//            winner is the lowest lane of a ballot over "my first-half maximum is the block maximum" (else of the
//            same ballot for the second half); its sign comes from a parallel ballot.  (Measured on B200: REDUX /
//            CREDUX reductions cost ~200 cycles each, the shuffle butterfly ~25 per step.)
//   iscale : __fdiv_rn(-127, max)            d : __fdiv_rn(1, iscale)
//   q      : min(127, nearest_int(__fmul_rn(iscale, x)))
// Returns d; w_lo / w_hi = the lane's two packed int8x4 words (qs words l and 32 + l); s_lo / s_hi = sums over the
// GROUP-lane group of the lane (GROUP = 4: the bsums of the 16-element groups l/4 and 8 + l/4; GROUP = 8: the pair
// sums bsums[2k] + bsums[2k+1] of 32 elements), valid in every lane of the group.  All-zero block: d = 0, q = 0
// and zero bsums (the reference leaves bsums unwritten there, :2568-2573).
template <int GROUP = 4>
__device__ __forceinline__ float quantize_block_q8_K(const float (&v)[8], int lane, uint32_t &w_lo, uint32_t &w_hi,
                                                     int &s_lo, int &s_hi) {
    (void)lane;
    // per-lane maxima of the two halves, first occurrence inside the lane (strict '>' in index order)
    uint32_t a_lo = 0u, a_hi = 0u;
    bool n_lo = false, n_hi = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t lo = __float_as_uint(fabsf(v[i])), hi = __float_as_uint(fabsf(v[4 + i]));
        const bool lo_ok = !(v[i] != v[i]), hi_ok = !(v[4 + i] != v[4 + i]);       // NaN never wins (amax < NaN is false)
        if (lo_ok && lo > a_lo) { a_lo = lo; n_lo = v[i] < 0.0f; }
        if (hi_ok && hi > a_hi) { a_hi = hi; n_hi = v[4 + i] < 0.0f; }
    }
    uint32_t abits = max(a_lo, a_hi);
    QT(16, abits);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) abits = max(abits, __shfl_xor_sync(0xffffffffu, abits, off));
    QT(17, abits);
    const unsigned in_lo = __ballot_sync(0xffffffffu, a_lo == abits), in_hi = __ballot_sync(0xffffffffu, a_hi == abits);
    const unsigned neg_lo = __ballot_sync(0xffffffffu, n_lo), neg_hi = __ballot_sync(0xffffffffu, n_hi);
    const bool neg = in_lo ? ((neg_lo >> (__ffs(in_lo) - 1)) & 1u) : ((neg_hi >> (__ffs(in_hi) - 1)) & 1u);
    QT(18, (unsigned)neg);
    int q[8];
    float d_out;
    if (abits == 0u) {
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = 0;
        d_out = 0.0f;
    } else {
        const float iscale = __fdiv_rn(-127.0f, __uint_as_float(abits | (neg ? 0x80000000u : 0u)));
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = min(127, nearest_int(__fmul_rn(iscale, v[i])));
        d_out = __fdiv_rn(1.0f, iscale);
    }
    w_lo = (uint32_t)(q[0] & 0xff) | ((uint32_t)(q[1] & 0xff) << 8) | ((uint32_t)(q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
    w_hi = (uint32_t)(q[4] & 0xff) | ((uint32_t)(q[5] & 0xff) << 8) | ((uint32_t)(q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
    QT(19, w_lo ^ w_hi);
    s_lo = q[0] + q[1] + q[2] + q[3];
    s_hi = q[4] + q[5] + q[6] + q[7];
#pragma unroll
    for (int off = 1; off < GROUP; off <<= 1) {
        s_lo += __shfl_xor_sync(0xffffffffu, s_lo, off);
        s_hi += __shfl_xor_sync(0xffffffffu, s_hi, off);
    }
    QT(20, (unsigned)(s_lo + s_hi) ^ __float_as_uint(d_out));
    return d_out;
}

}  // namespace q4k
