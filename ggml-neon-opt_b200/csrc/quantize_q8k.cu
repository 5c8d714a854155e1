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
    // programmatic dependent launch on both sides: the decode kernel behind this one may start its prologue (barriers, the
    // first weight tiles — they do not depend on the activations) right away; this kernel waits for the one in front of it
    // before it reads x or overwrites the workspace that kernel may still be reading
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int lane = threadIdx.x & 31;
    const int64_t blk = (int64_t)blockIdx.x * QUANT_WARPS + (threadIdx.x >> 5);
    if (blk >= nblocks_total) return;
    const int64_t row = blk / nb_per_row, col = blk - row * nb_per_row;
    const float *src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + row * x_row_stride_bytes) + col * QK_K;
    uint8_t *dst = y + blk * BLK_Q8K;   // rows are packed: row stride = nb_per_row * 292

    // plain (coherent) loads, not ld.global.nc: under programmatic dependent launch this kernel's lifetime overlaps the kernel
    // that produces x (peer GPUs storing over NVLink included)
    const float4 a = ld_act16(src + 4 * lane);
    const float4 b = ld_act16(src + 128 + 4 * lane);
    const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};

    uint32_t w_lo, w_hi;
    int s_lo, s_hi;
    const float d_out = quantize_block_q8_K(v, lane, w_lo, w_hi, s_lo, s_hi);
    uint32_t *qs = reinterpret_cast<uint32_t *>(dst + Q8K_QS);
    qs[lane]      = w_lo;
    qs[32 + lane] = w_hi;
    int16_t *bsums = reinterpret_cast<int16_t *>(dst + Q8K_BSUMS);
    if ((lane & 3) == 0) {
        bsums[lane >> 2]       = (int16_t)s_lo;     // group (4l)/16 = l/4
        bsums[8 + (lane >> 2)] = (int16_t)s_hi;     // group 8 + l/4
    }
    if (lane == 0) *reinterpret_cast<float *>(dst) = d_out;
}

cudaError_t launch_quantize_q8_K(const float *x, void *y, int64_t k, int64_t nrows, int64_t x_row_stride_bytes,
                                 cudaStream_t stream) {
    const int64_t nb_per_row = k / QK_K;
    const int64_t total = nb_per_row * nrows;
    if (total == 0) return cudaSuccess;
    const int64_t grid = (total + QUANT_WARPS - 1) / QUANT_WARPS;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(QUANT_WARPS * 32);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = knobs().pdl ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, quantize_q8_K_kernel, x, (uint8_t *)y, nb_per_row, total, x_row_stride_bytes);
    count_launch();
    return e;
}

}  // namespace q4k
