// quantize_q8k.cu — device activation quantizer, bit-exact with the reference's quantize_row_q8_K
// (L/ggml/src/ggml-quants.c:2555-2592; every arch forwards to it: arch/x86/quants.c:493-495).
//
// One warp per 256-float block.  Lane l owns floats [4l, 4l+4) and [128+4l, 128+4l+4): two coalesced
// 16-byte loads per lane, two coalesced 4-byte stores of packed int8 per lane.
//   absmax      : per-lane scan in index order (strict '>'), then a shuffle butterfly that prefers the
//                 larger |x| and, on ties, the smaller index == "first occurrence wins" of the scalar loop
//   iscale      : __fdiv_rn(-127, max)           one IEEE divide (the reference's divss)
//   q           : nearest_int(__fmul_rn(iscale, x)), min 127      two roundings, never an FMA
//   bsums       : 4 values per lane + butterfly over 4 adjacent lanes (integers: order-free)
//   d           : __fdiv_rn(1, iscale)
// HBM-bound: reads 1024 B, writes 292 B per block.  This TU is also compiled with -fmad=false.
#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

constexpr int QUANT_WARPS = 8;

__global__ void __launch_bounds__(QUANT_WARPS * 32)
quantize_q8_K_kernel(const float *__restrict__ x, uint8_t *__restrict__ y, int64_t nb_per_row,
                     int64_t nblocks_total, int64_t x_row_stride_bytes) {
    const int lane = threadIdx.x & 31;
    const int64_t blk = (int64_t)blockIdx.x * QUANT_WARPS + (threadIdx.x >> 5);
    if (blk >= nblocks_total) return;
    const int64_t row = blk / nb_per_row, col = blk - row * nb_per_row;
    const float *src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + row * x_row_stride_bytes) + col * QK_K;
    uint8_t *dst = y + blk * BLK_Q8K;   // rows are packed: row stride = nb_per_row * 292

    const float4 a = __ldg(reinterpret_cast<const float4 *>(src) + lane);
    const float4 b = __ldg(reinterpret_cast<const float4 *>(src) + 32 + lane);
    const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};

    // per-lane first-occurrence absmax over indices 4l..4l+3 then 128+4l..128+4l+3 (ascending)
    float amax = 0.0f, smax = 0.0f;
    int idx = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float ax = fabsf(v[i]);
        if (ax > amax) { amax = ax; smax = v[i]; idx = (i < 4 ? 4 * lane + i : 128 + 4 * lane + (i - 4)); }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, amax, off);
        const float os = __shfl_xor_sync(0xffffffffu, smax, off);
        const int   oi = __shfl_xor_sync(0xffffffffu, idx, off);
        if (oa > amax || (oa == amax && oi < idx)) { amax = oa; smax = os; idx = oi; }
    }

    int q[8];
    float d_out;
    if (amax == 0.0f) {                       // :2568-2573 (we also zero the bsums, see q4k_b200.h)
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = 0;
        d_out = 0.0f;
    } else {
        const float iscale = __fdiv_rn(-127.0f, smax);
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = min(127, nearest_int(__fmul_rn(iscale, v[i])));
        d_out = __fdiv_rn(1.0f, iscale);
    }

    auto pack4 = [](int q0, int q1, int q2, int q3) -> uint32_t {
        return (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) | ((uint32_t)(q3 & 0xff) << 24);
    };
    uint32_t *qs = reinterpret_cast<uint32_t *>(dst + Q8K_QS);
    qs[lane]      = pack4(q[0], q[1], q[2], q[3]);
    qs[32 + lane] = pack4(q[4], q[5], q[6], q[7]);

    int s_lo = q[0] + q[1] + q[2] + q[3];     // group (4l)/16     = l/4
    int s_hi = q[4] + q[5] + q[6] + q[7];     // group 8 + l/4
    s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 1);
    s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 1);
    s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 2);
    s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 2);
    int16_t *bsums = reinterpret_cast<int16_t *>(dst + Q8K_BSUMS);
    if ((lane & 3) == 0) {
        bsums[lane >> 2]       = (int16_t)s_lo;
        bsums[8 + (lane >> 2)] = (int16_t)s_hi;
    }
    if (lane == 0) *reinterpret_cast<float *>(dst) = d_out;
}

cudaError_t launch_quantize_q8_K(const float *x, void *y, int64_t k, int64_t nrows, int64_t x_row_stride_bytes,
                                 cudaStream_t stream) {
    const int64_t nb_per_row = k / QK_K;
    const int64_t total = nb_per_row * nrows;
    if (total == 0) return cudaSuccess;
    const int64_t grid = (total + QUANT_WARPS - 1) / QUANT_WARPS;
    quantize_q8_K_kernel<<<(unsigned)grid, QUANT_WARPS * 32, 0, stream>>>(x, (uint8_t *)y, nb_per_row, total, x_row_stride_bytes);
    count_launch();
    return cudaGetLastError();
}

}  // namespace q4k
