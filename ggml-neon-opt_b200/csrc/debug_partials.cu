// debug_partials.cu — dumps the integer stage of the dot product:
//   p[j]  = sum_{l<32} q4[32j+l] * q8[32j+l]        j = 0..7      int32
//   sumi  = sum_j sc[j] * p[j]                                     int32
//   summ  = sum_j m[j] * (bsums[2j] + bsums[2j+1])                 int32
// for a range of weight rows against every activation row.  Parity tooling for the "bit-exact
// per-sub-block int32 partial sums" bar; it shares the header/nibble/scale unpack helpers and the
// dp4a formulation with the production GEMV (q4k_common.cuh), one octet of lanes per superblock.
#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

__global__ void __launch_bounds__(256)
debug_partials_kernel(const uint8_t *__restrict__ w, int nb, int64_t w_stride, int64_t row0, int64_t nrows,
                      const uint8_t *__restrict__ xq, int64_t N, int64_t xq_stride,
                      int32_t *__restrict__ p, int32_t *__restrict__ sumi, int32_t *__restrict__ summ) {
    const int c = threadIdx.x & 7;
    const int64_t total = N * nrows * nb;
    const int64_t raw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;    // (n, r, b)
    const bool valid = raw < total;                 // no early exit: full-mask shuffles below
    const int64_t item = valid ? raw : total - 1;
    const int b = (int)(item % nb);
    const int64_t r = (item / nb) % nrows, n = item / ((int64_t)nb * nrows);

    const uint8_t *blk = w + (row0 + r) * w_stride + (int64_t)b * BLK_Q4K;
    const uint4 hdr = ldg_stream16(blk);
    const uint4 q = ldg_stream16(blk + 16 + 16 * c);
    const uint8_t *yb = xq + n * xq_stride + (int64_t)b * BLK_Q8K;
    const int g = c >> 1, h = c & 1;

    const uint4 sm = unpack_scales_mins(hdr);
    const uint32_t scp = byte_pair(sm.x, sm.y, g), mp = byte_pair(sm.z, sm.w, g);
    const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
    const uint32_t *ylo = reinterpret_cast<const uint32_t *>(yb + Q8K_QS + 64 * g + 16 * h);
    const uint32_t *yhi = ylo + 8;
    int pl = 0, ph = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        pl = __dp4a((int)(qw[i] & 0x0f0f0f0fu), (int)ylo[i], pl);
        ph = __dp4a((int)((qw[i] >> 4) & 0x0f0f0f0fu), (int)yhi[i], ph);
    }
    // the two halves (h = 0, 1) of a sub-block live in adjacent lanes
    pl += __shfl_xor_sync(0xffffffffu, pl, 1);
    ph += __shfl_xor_sync(0xffffffffu, ph, 1);
    int si = 0, sm_i = 0;
    if (h == 0 && valid) {
        const int16_t *bs = reinterpret_cast<const int16_t *>(yb + Q8K_BSUMS);
        si = (int)(scp & 0xff) * pl + (int)(scp >> 8) * ph;
        sm_i = (int)(mp & 0xff) * ((int)bs[4 * g] + (int)bs[4 * g + 1]) + (int)(mp >> 8) * ((int)bs[4 * g + 2] + (int)bs[4 * g + 3]);
        p[item * 8 + 2 * g] = pl;
        p[item * 8 + 2 * g + 1] = ph;
    }
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) {
        si += __shfl_xor_sync(0xffffffffu, si, off);
        sm_i += __shfl_xor_sync(0xffffffffu, sm_i, off);
    }
    if (c == 0 && valid) { sumi[item] = si; summ[item] = sm_i; }
}

cudaError_t launch_debug_partials(const void *w, int64_t K, int64_t w_row_stride_bytes, int64_t row0, int64_t nrows,
                                  const void *xq, int64_t N, int64_t xq_row_stride_bytes,
                                  int32_t *p, int32_t *sumi, int32_t *summ, cudaStream_t stream) {
    const int nb = (int)(K / QK_K);
    const int64_t threads = N * nrows * nb * 8;
    if (threads == 0) return cudaSuccess;
    const int64_t grid = (threads + 255) / 256;
    debug_partials_kernel<<<(unsigned)grid, 256, 0, stream>>>((const uint8_t *)w, nb, w_row_stride_bytes, row0, nrows,
                                                              (const uint8_t *)xq, N, xq_row_stride_bytes, p, sumi, summ);
    count_launch();
    return cudaGetLastError();
}

}  // namespace q4k
