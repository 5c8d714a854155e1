// gemv_q4k.cu — decode GEMV (batch 1..8): dst[n][m] = vec_dot(W row m, Q8_K(x row n)).
// Replaces the chunked vec_dot loop of ggml_compute_forward_mul_mat_one_chunk
// (L/ggml/src/ggml-cpu/ggml-cpu.c:1115-1203) + ggml_vec_dot_q4_K_q8_K (arch/x86/quants.c:1741-1823).
//
// HBM-bound by design: every weight byte is read exactly once with 16-byte streaming loads, the
// (small) Q8_K activations are re-laid out into shared memory once per CTA.  No tensor cores.
//
// Mapping.  A block_q4_K is nine 16 B chunks: chunk 0 = {d, dmin, 12 scale bytes}, chunks 1..8 = nibbles.
// Eight lanes (an "octet") own one weight row; lane c of the octet owns nibble chunk c of the current
// superblock: group g = c/2 (64 elements), half h = c%2, i.e. low nibbles = elements 64g+16h..+15 of
// sub-block 2g and high nibbles = elements 64g+32+16h..+15 of sub-block 2g+1.  A warp therefore streams
// four rows (x R row sets) and walks the superblocks of those rows in order, U at a time.
//
// Float order.  With EXACT the kernel reproduces the reference's AVX2 accumulation bit for bit
// (arch/x86/quants.c:1760-1823): 32-bit lane t of the AVX2 `sumi` is word t%4 of half t/4 summed over
// the four groups, so a 3-shuffle integer transpose-reduce over the four lanes that share h leaves each
// of the eight lanes of the octet holding exactly one AVX2 lane sum; it then runs that lane's FMA chain
// acc = fma(d, (float)sumi_t, acc) over the superblocks in order.  The mins use four more chains
// (acc_m, lanes 0..3), and the final horizontal sums follow hsum_float_8 / movehl / movehdup order.
// Without EXACT each lane folds its 2x16 elements into one int and the eight lanes are summed at the
// end (same integers, different float association; within 1e-5 of the reference, see tests).
#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

struct GemvParams {
    const uint8_t *w;
    int64_t M, w_stride;
    int nb;                 // superblocks per row
    const uint8_t *xq;
    int64_t xq_stride;
    float *dst;
    int64_t dst_stride;     // floats between dst columns
    int kt;                 // superblocks per shared-memory K tile
    int n_ktiles;
    int64_t n_row_tiles;
};

// Q8_K -> shared memory, kernel-friendly layout: qs 16 B aligned, pair-summed bsums, scales apart.
template <int N>
__device__ __forceinline__ void stage_q8(const GemvParams &p, int kb0, int nkb, uint32_t *qs_s, int16_t *ps_s, float *d_s) {
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < N * nkb * 64; i += nt) {
        const int n = i / (nkb * 64), r = i - n * nkb * 64, b = r >> 6, wd = r & 63;
        const uint8_t *blk = p.xq + n * p.xq_stride + (int64_t)(kb0 + b) * BLK_Q8K;
        qs_s[(n * p.kt + b) * 64 + wd] = __ldg(reinterpret_cast<const uint32_t *>(blk + Q8K_QS) + wd);
    }
    for (int i = tid; i < N * nkb * 8; i += nt) {
        const int n = i / (nkb * 8), r = i - n * nkb * 8, b = r >> 3, k = r & 7;
        const uint8_t *blk = p.xq + n * p.xq_stride + (int64_t)(kb0 + b) * BLK_Q8K;
        const int16_t *bs = reinterpret_cast<const int16_t *>(blk + Q8K_BSUMS);
        ps_s[(n * p.kt + b) * 8 + k] = (int16_t)(bs[2 * k] + bs[2 * k + 1]);   // _mm_hadd_epi16 of the bsums
    }
    for (int i = tid; i < N * nkb; i += nt) {
        const int n = i / nkb, b = i - n * nkb;
        d_s[n * p.kt + b] = *reinterpret_cast<const float *>(p.xq + n * p.xq_stride + (int64_t)(kb0 + b) * BLK_Q8K);
    }
}

// one superblock of one row against all N columns
template <int N, bool EXACT>
__device__ __forceinline__ void block_dot(const uint4 hdr, const uint4 q, int bb, int kt, int c,
                                          const uint32_t *qs_s, const int16_t *ps_s, const float *d_s,
                                          float (&acc)[N], float (&accm)[N]) {
    const int g = c >> 1, h = c & 1, u4 = c & 3;
    const uint4 sm = unpack_scales_mins(hdr);
    const uint32_t scp = byte_pair(sm.x, sm.y, g);
    const int sc_lo = scp & 0xff, sc_hi = scp >> 8;
    const int mpair = (int)byte_pair(sm.z, sm.w, u4);
    const float xd = header_d(hdr), xdm = header_dmin(hdr);
    const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
    int lo[4], hi[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        lo[w] = (int)(qw[w] & 0x0f0f0f0fu);
        hi[w] = (int)((qw[w] >> 4) & 0x0f0f0f0fu);
    }
#pragma unroll
    for (int n = 0; n < N; ++n) {
        const int blk = n * kt + bb;
        const uint4 ql = *reinterpret_cast<const uint4 *>(qs_s + blk * 64 + 16 * g + 4 * h);
        const uint4 qh = *reinterpret_cast<const uint4 *>(qs_s + blk * 64 + 16 * g + 8 + 4 * h);
        const float yd = d_s[blk];
        const float dall = __fmul_rn(yd, xd);
        const float dmn = __fmul_rn(-yd, xdm);
        const int pp = *reinterpret_cast<const int *>(ps_s + blk * 8 + 2 * u4);
        accm[n] = __fmaf_rn(dmn, (float)__dp2a_lo(pp, mpair, 0), accm[n]);
        const int y_lo[4] = {(int)ql.x, (int)ql.y, (int)ql.z, (int)ql.w};
        const int y_hi[4] = {(int)qh.x, (int)qh.y, (int)qh.z, (int)qh.w};
        if (EXACT) {
            int v[4];
#pragma unroll
            for (int w = 0; w < 4; ++w) v[w] = __dp4a(lo[w], y_lo[w], 0) * sc_lo + __dp4a(hi[w], y_hi[w], 0) * sc_hi;
            const bool g0 = c & 2, g1 = c & 4;
            int ka = g0 ? v[2] : v[0], kb = g0 ? v[3] : v[1];
            const int sa = g0 ? v[0] : v[2], sb = g0 ? v[1] : v[3];
            ka += __shfl_xor_sync(0xffffffffu, sa, 2);
            kb += __shfl_xor_sync(0xffffffffu, sb, 2);
            int keep = g1 ? kb : ka;
            const int send = g1 ? ka : kb;
            keep += __shfl_xor_sync(0xffffffffu, send, 4);
            acc[n] = __fmaf_rn(dall, (float)keep, acc[n]);     // AVX2 lane t = 4h + 2*g0 + g1
        } else {
            int pl = 0, ph = 0;
#pragma unroll
            for (int w = 0; w < 4; ++w) { pl = __dp4a(lo[w], y_lo[w], pl); ph = __dp4a(hi[w], y_hi[w], ph); }
            acc[n] = __fmaf_rn(dall, (float)(pl * sc_lo + ph * sc_hi), acc[n]);
        }
    }
}

// THREADS/32 warps per CTA, each warp streams 4*R rows; U superblocks of every row are in flight at once.
template <int N, bool EXACT, int R, int U, int THREADS>
__global__ void __launch_bounds__(THREADS)
gemv_q4k_q8k_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t *qs_s = reinterpret_cast<uint32_t *>(smem);
    int16_t *ps_s = reinterpret_cast<int16_t *>(smem + (size_t)N * p.kt * 256);
    float *d_s = reinterpret_cast<float *>(smem + (size_t)N * p.kt * (256 + 16));

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int o = lane >> 3, c = lane & 7;
    constexpr int ROWS_PER_CTA = (THREADS / 32) * 4 * R;

    if (p.n_ktiles == 1) {                       // activations fit: stage once, stay resident for all row tiles
        stage_q8<N>(p, 0, p.nb, qs_s, ps_s, d_s);
        __syncthreads();
    }

    for (int64_t tile = blockIdx.x; tile < p.n_row_tiles; tile += gridDim.x) {
        int64_t row[R];
        const uint8_t *wrow[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            row[r] = tile * ROWS_PER_CTA + warp * 4 * R + r * 4 + o;
            const int64_t rc = row[r] < p.M ? row[r] : p.M - 1;          // clamp: loads stay in bounds, store is skipped
            wrow[r] = p.w + rc * p.w_stride + 16 * c;
        }
        float acc[R][N], accm[R][N];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int n = 0; n < N; ++n) { acc[r][n] = 0.0f; accm[r][n] = 0.0f; }

        for (int ktile = 0; ktile < p.n_ktiles; ++ktile) {
            const int kb0 = ktile * p.kt;
            const int nkb = min(p.kt, p.nb - kb0);
            if (p.n_ktiles > 1) {
                __syncthreads();
                stage_q8<N>(p, kb0, nkb, qs_s, ps_s, d_s);
                __syncthreads();
            }
            for (int b = 0; b < nkb; b += U) {
                uint4 hdr[U][R], q[U][R];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int bb = min(b + u, nkb - 1);                    // tail: re-read the last block, result unused
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const uint8_t *blk = wrow[r] + (int64_t)(kb0 + bb) * BLK_Q4K;
                        hdr[u][r] = ldg_stream16(blk - 16 * c);
                        q[u][r] = ldg_stream16(blk + 16);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (b + u < nkb) {
#pragma unroll
                        for (int r = 0; r < R; ++r)
                            block_dot<N, EXACT>(hdr[u][r], q[u][r], b + u, p.kt, c, qs_s, ps_s, d_s, acc[r], accm[r]);
                    }
                }
            }
        }

#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int n = 0; n < N; ++n) {
                float a = acc[r][n], m = accm[r][n];
                a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 1));     // hsum_float_8: lanes t and t+4
                a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 2));     //   (r0+r2), (r1+r3)
                a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 4));     //   sum
                m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 2));     // acc_m + movehl(acc_m)
                m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 1));     // + movehdup
                if (c == 0 && row[r] < p.M) p.dst[(int64_t)n * p.dst_stride + row[r]] = __fadd_rn(a, m);
            }
    }
}

static size_t gemv_smem_bytes(int N, int kt) { return (size_t)N * kt * (256 + 16 + 4); }

template <int N, bool EXACT, int R, int U, int THREADS>
static cudaError_t launch_cfg(GemvParams p, cudaStream_t stream) {
    auto kern = gemv_q4k_q8k_kernel<N, EXACT, R, U, THREADS>;
    constexpr size_t SMEM_BUDGET = 96 * 1024;
    const int64_t kt_fit = (int64_t)(SMEM_BUDGET / ((size_t)N * 276));
    p.kt = (int)(p.nb < kt_fit ? p.nb : kt_fit);
    p.n_ktiles = (p.nb + p.kt - 1) / p.kt;
    const size_t smem = gemv_smem_bytes(N, p.kt);
    if (const cudaError_t ea = ensure_max_dynamic_smem((const void *)kern, (int)SMEM_BUDGET)) return ea;
    constexpr int ROWS_PER_CTA = (THREADS / 32) * 4 * R;
    p.n_row_tiles = (p.M + ROWS_PER_CTA - 1) / ROWS_PER_CTA;
    int per_sm = 1;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, THREADS, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    const int64_t resident = (int64_t)sm_count() * per_sm;
    const int64_t grid = p.n_row_tiles < resident ? p.n_row_tiles : resident;
    kern<<<(unsigned)grid, THREADS, smem, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}


template <int N, bool EXACT>
static cudaError_t launch_n(const GemvParams &p, cudaStream_t stream) {
    if (N > 4) return launch_cfg<N, EXACT, 1, 2, 128>(p, stream);
    switch (knobs().gemv_config) {
        case 1: return launch_cfg<N, EXACT, 1, 4, 128>(p, stream);
        case 2: return launch_cfg<N, EXACT, 2, 2, 256>(p, stream);
        case 3: return launch_cfg<N, EXACT, 1, 8, 128>(p, stream);
        case 4: return launch_cfg<N, EXACT, 2, 4, 256>(p, stream);
        case 5: return launch_cfg<N, EXACT, 1, 4, 256>(p, stream);
        case 6: return launch_cfg<N, EXACT, 2, 4, 128>(p, stream);
        default: break;
    }
    // small matrices: small CTAs so every SM gets several; large ones: 64-row CTAs, two row sets per lane
    const int64_t tiles16 = (p.M + 15) / 16;
    if (tiles16 <= (int64_t)sm_count() * 8) return launch_cfg<N, EXACT, 1, 4, 128>(p, stream);
    return launch_cfg<N, EXACT, 2, 2, 256>(p, stream);
}

template <bool EXACT>
static cudaError_t launch_e(int N, const GemvParams &p, cudaStream_t stream) {
    // Only N == 1 is instantiated: this direct-load kernel is the last resort for ONE activation row too long for a shared-memory
    // residency (K > ~200 000); wider batches are halved by the caller until they fit the staged kernels or reach one column.
    // (The other seven batch sizes were 98 more kernels in the library for no reachable launch.)
    if (N == 1) return launch_n<1, EXACT>(p, stream);
    return cudaErrorNotSupported;
}

cudaError_t launch_gemv_q4k_q8k(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                const void *xq, int N, int64_t xq_row_stride_bytes,
                                float *dst, int64_t dst_col_stride_bytes, bool exact_order, cudaStream_t stream) {
    if (M == 0 || N == 0) return cudaSuccess;
    const int g_gemv_config = knobs().gemv_config;
    if (g_gemv_config == 0 || g_gemv_config >= 10) {      // production path: TMA-bulk staged kernel
        const cudaError_t e = launch_gemv_q4k_q8k_tma(w, M, K, w_row_stride_bytes, xq, nullptr, N, xq_row_stride_bytes, dst,
                                                      dst_col_stride_bytes, exact_order,
                                                      g_gemv_config >= 10 ? g_gemv_config - 10 : 0, knobs().pdl != 0, stream);
        if (e != cudaErrorNotSupported) return e;
        // activations too large for one shared-memory residency (8 columns of K = 28672 ...): half the columns per launch —
        // the weights stream twice, which still beats the K-tiled direct-load kernel below by an order of magnitude
        if (N > 1) {
            const int n0 = N / 2;
            const cudaError_t e0 = launch_gemv_q4k_q8k(w, M, K, w_row_stride_bytes, xq, n0, xq_row_stride_bytes, dst, dst_col_stride_bytes,
                                                       exact_order, stream);
            if (e0 != cudaSuccess) return e0;
            return launch_gemv_q4k_q8k(w, M, K, w_row_stride_bytes, (const uint8_t *)xq + (size_t)n0 * xq_row_stride_bytes, N - n0,
                                       xq_row_stride_bytes, dst + (size_t)n0 * (dst_col_stride_bytes / 4), dst_col_stride_bytes, exact_order, stream);
        }
    }
    GemvParams p;
    p.w = (const uint8_t *)w; p.M = M; p.w_stride = w_row_stride_bytes; p.nb = (int)(K / QK_K);
    p.xq = (const uint8_t *)xq; p.xq_stride = xq_row_stride_bytes;
    p.dst = dst; p.dst_stride = dst_col_stride_bytes / 4;
    p.kt = 0; p.n_ktiles = 0; p.n_row_tiles = 0;
    return exact_order ? launch_e<true>(N, p, stream) : launch_e<false>(N, p, stream);
}

}  // namespace q4k
