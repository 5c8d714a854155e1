// gemv_q4k_tma.cu — decode GEMV (batch 1..8), TMA-bulk staged.  dst[n][m] = vec_dot(W row m, Q8_K(x row n)).
// Replaces ggml_compute_forward_mul_mat_one_chunk's vec_dot loop (L/ggml/src/ggml-cpu/ggml-cpu.c:1115-1203)
// + ggml_vec_dot_q4_K_q8_K (L/ggml/src/ggml-cpu/arch/x86/quants.c:1741-1823) for N <= 8.
//
// HBM-bound by design; no tensor cores.  Every weight byte crosses HBM once, moved by the copy engine:
//   * each warp owns a private ring of S stages; a stage holds a 4-row x 8-superblock tile of block_q4_K
//     (4 x 1152 contiguous bytes) fetched by lane 0 with cp.async.bulk (UBLKCP) + mbarrier complete_tx.
//     S-1 tiles (~4.6 KB each) are in flight per warp with no registers tied up.
//   * one lane per superblock ("lane (r, b)": row r of the tile, superblock b): the 16 B header is unpacked
//     once per superblock (no redundancy across lanes), nibbles come out of shared memory with conflict-free
//     LDS.128 (144 B block stride = 36 words = 4 banks per lane step).
//   * the Q8_K activations are re-laid out once per CTA: 272 B per block (256 qs + the 8 pair-summed bsums),
//     68 words = 4 banks per block step, so eight lanes on eight different blocks never conflict and lanes
//     on the same block broadcast.
//   * rows are dealt to CTAs in balanced contiguous ranges; warps of a CTA pull 4-row groups from a
//     shared-memory counter, so the tail inside an SM is one row group.
//
// Float order.  EXACT reproduces the reference's AVX2 accumulation bit for bit: the integer lane sums
// sumi_t (t = 0..7: word t of every 32-byte nibble group) and the four min products of every superblock
// are parked in a per-warp scratch, then lane (r, t) runs the reference's FMA chain over the superblocks
// in order: acc_t = fma(d, (float)sumi_t, acc_t), acc_m = fma(dmin, (float)prod, acc_m); the final
// horizontal adds follow hsum_float_8 / movehl / movehdup.  Without EXACT every lane folds its superblock
// into one int and lanes are summed at the end (same integers; fp32 within 1e-5 of the reference).
//
// The high nibbles are never shifted: dp4a.u32.s32 on (w & 0xf0f0f0f0) gives 16 x the high-nibble dot
// product, the low side is multiplied by 16 x scale, and the (exact) >> 4 happens once per lane sum.
#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

constexpr int TILE_ROWS = 4;
constexpr int TILE_SB = 8;
constexpr int ROW_BYTES = TILE_SB * BLK_Q4K;        // 1152
constexpr int STAGE_BYTES = TILE_ROWS * ROW_BYTES;  // 4608
constexpr int Q8_STRIDE = 272;                      // 256 qs + 8 x int16 pair sums
constexpr int SCR_WORDS = 32 * 20 + 32;             // per column scratch: 20 words per (r, b) + row skew
constexpr int WARPS = 8;

struct TmaGemvParams {
    const uint8_t *w;
    int64_t M, w_stride;
    int nb;
    const uint8_t *xq;
    int64_t xq_stride;
    float *dst;
    int64_t dst_stride;
    int64_t n_groups;       // ceil(M / 4)
    int stages;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ int dp4a_u8s8(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

}  // namespace

namespace {

__device__ __forceinline__ uint4 lds128(const uint8_t *p) { return *reinterpret_cast<const uint4 *>(p); }

template <int N>
__device__ __forceinline__ void stage_activations(const TmaGemvParams &p, uint8_t *q8_s, float *d_s) {
    const int tid = threadIdx.x, nt = blockDim.x, nb = p.nb;
    for (int i = tid; i < N * nb * 64; i += nt) {
        const int blk = i >> 6, wd = i & 63, n = blk / nb, b = blk - n * nb;
        const uint8_t *src = p.xq + n * p.xq_stride + (int64_t)b * BLK_Q8K;
        reinterpret_cast<uint32_t *>(q8_s + (size_t)blk * Q8_STRIDE)[wd] = __ldg(reinterpret_cast<const uint32_t *>(src + Q8K_QS) + wd);
    }
    for (int i = tid; i < N * nb * 8; i += nt) {
        const int blk = i >> 3, k = i & 7, n = blk / nb, b = blk - n * nb;
        const int16_t *bs = reinterpret_cast<const int16_t *>(p.xq + n * p.xq_stride + (int64_t)b * BLK_Q8K + Q8K_BSUMS);
        reinterpret_cast<int16_t *>(q8_s + (size_t)blk * Q8_STRIDE + 256)[k] = (int16_t)(bs[2 * k] + bs[2 * k + 1]);   // hadd_epi16
    }
    for (int i = tid; i < N * nb; i += nt) {
        const int n = i / nb, b = i - n * nb;
        d_s[i] = *reinterpret_cast<const float *>(p.xq + n * p.xq_stride + (int64_t)b * BLK_Q8K);
    }
}

template <int N, bool EXACT>
__global__ void __launch_bounds__(WARPS * 32)
gemv_q4k_tma_kernel(const TmaGemvParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int nb = p.nb, S = p.stages;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r = lane >> 3, b = lane & 7;        // int phase: row r of the tile, superblock b; chain phase: (r, t = b)

    size_t off = 0;
    uint8_t *q8_s = smem;                                  off += (size_t)N * nb * Q8_STRIDE;
    float *d_s = reinterpret_cast<float *>(smem + off);    off += ((size_t)N * nb * 4 + 127) & ~(size_t)127;
    off = (off + 127) & ~(size_t)127;
    uint8_t *wbuf = smem + off + (size_t)warp * S * STAGE_BYTES;   off += (size_t)WARPS * S * STAGE_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + off) + warp * S;   off += (size_t)WARPS * S * 8;
    int2 *meta = reinterpret_cast<int2 *>(smem + off) + warp * S;   off += (size_t)WARPS * S * 8;
    int *counter = reinterpret_cast<int *>(smem + off);    off += 16;
    int *scratch = reinterpret_cast<int *>(smem + off) + (EXACT ? warp * N * SCR_WORDS : 0);

    // balanced contiguous range of 4-row groups for this CTA
    const int64_t g0 = p.n_groups * blockIdx.x / gridDim.x, g1 = p.n_groups * (blockIdx.x + 1) / gridDim.x;
    const int ksteps = (nb + TILE_SB - 1) / TILE_SB;

    if (threadIdx.x == 0) *counter = 0;
    if (lane == 0) {
        for (int s = 0; s < S; ++s) mbar_init(smem_u32(&bars[s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    stage_activations<N>(p, q8_s, d_s);
    __syncthreads();

    // ---- producer state (warp-uniform) ----
    int64_t cur_rg = -1;
    int prod_ks = ksteps;       // forces a fetch on first use
    bool exhausted = false;
    auto produce = [&](int s) {
        if (!exhausted && prod_ks == ksteps) {
            int nxt = 0;
            if (lane == 0) nxt = atomicAdd(counter, 1);
            nxt = __shfl_sync(0xffffffffu, nxt, 0);
            cur_rg = g0 + nxt;
            prod_ks = 0;
            if (cur_rg >= g1) exhausted = true;
        }
        const uint32_t bar = smem_u32(&bars[s]);
        if (exhausted) {
            if (lane == 0) {
                meta[s] = make_int2(-1, 0);
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
            }
            return;
        }
        const int nsb = min(TILE_SB, nb - prod_ks * TILE_SB);
        const uint32_t bytes = (uint32_t)nsb * BLK_Q4K;
        if (lane == 0) {
            meta[s] = make_int2((int)cur_rg, prod_ks);
            fence_proxy_async();                      // our generic reads of this stage are done (after __syncwarp)
            mbar_expect_tx(bar, bytes * TILE_ROWS);
        }
        __syncwarp();
        if (lane < TILE_ROWS) {
            int64_t row = cur_rg * TILE_ROWS + lane;
            if (row >= p.M) row = p.M - 1;            // ragged last group: stay in bounds, result is not stored
            bulk_g2s(smem_u32(wbuf + (size_t)s * STAGE_BYTES + lane * ROW_BYTES),
                     p.w + row * p.w_stride + (int64_t)prod_ks * ROW_BYTES, bytes, bar);
        }
        ++prod_ks;
    };

    for (int s = 0; s < S; ++s) produce(s);

    float acc[N], accm[N];
#pragma unroll
    for (int n = 0; n < N; ++n) { acc[n] = 0.0f; accm[n] = 0.0f; }

    for (int it = 0;; ++it) {
        const int s = it % S;
        mbar_wait(smem_u32(&bars[s]), (uint32_t)((it / S) & 1));
        const int2 mt = meta[s];
        if (mt.x < 0) break;
        const int ks = mt.y;
        const int nsb = min(TILE_SB, nb - ks * TILE_SB);
        const uint8_t *wst = wbuf + (size_t)s * STAGE_BYTES;

        if (b < nsb) {
            const uint8_t *blk = wst + r * ROW_BYTES + b * BLK_Q4K;
            const int kb = ks * TILE_SB + b;
            const uint4 hdr = lds128(blk);
            const uint4 sm = unpack_scales_mins(hdr);
            const float xd = header_d(hdr), xdm = header_dmin(hdr);
            int sum16[N][EXACT ? 8 : 1];
#pragma unroll
            for (int n = 0; n < N; ++n)
#pragma unroll
                for (int t = 0; t < (EXACT ? 8 : 1); ++t) sum16[n][t] = 0;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const uint4 qa = lds128(blk + 16 + 32 * g), qb = lds128(blk + 32 + 32 * g);
                const uint32_t q[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
                const uint32_t scp = byte_pair(sm.x, sm.y, g);
                const int sc_lo16 = (int)((scp & 0xffu) << 4), sc_hi = (int)(scp >> 8);
#pragma unroll
                for (int n = 0; n < N; ++n) {
                    const uint8_t *yb = q8_s + (size_t)(n * nb + kb) * Q8_STRIDE + 64 * g;
                    const uint4 y0 = lds128(yb), y1 = lds128(yb + 16), y2 = lds128(yb + 32), y3 = lds128(yb + 48);
                    const uint32_t ylo[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
                    const uint32_t yhi[8] = {y2.x, y2.y, y2.z, y2.w, y3.x, y3.y, y3.z, y3.w};
                    if (EXACT) {
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            const int dl = __dp4a((int)(q[t] & 0x0f0f0f0fu), (int)ylo[t], 0);
                            const int dh = dp4a_u8s8(q[t] & 0xf0f0f0f0u, yhi[t], 0);
                            sum16[n][t] += sc_lo16 * dl + sc_hi * dh;
                        }
                    } else {
                        int pl = 0, ph = 0;
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            pl = __dp4a((int)(q[t] & 0x0f0f0f0fu), (int)ylo[t], pl);
                            ph = dp4a_u8s8(q[t] & 0xf0f0f0f0u, yhi[t], ph);
                        }
                        sum16[n][0] += sc_lo16 * pl + sc_hi * ph;
                    }
                }
            }
#pragma unroll
            for (int n = 0; n < N; ++n) {
                const uint8_t *yblk = q8_s + (size_t)(n * nb + kb) * Q8_STRIDE;
                const uint4 ps = lds128(yblk + 256);
                const float yd = d_s[n * nb + kb];
                const float dall = __fmul_rn(yd, xd), dmn = __fmul_rn(-yd, xdm);
                const int p0 = __dp2a_lo((int)ps.x, (int)byte_pair(sm.z, sm.w, 0), 0);
                const int p1 = __dp2a_lo((int)ps.y, (int)byte_pair(sm.z, sm.w, 1), 0);
                const int p2 = __dp2a_lo((int)ps.z, (int)byte_pair(sm.z, sm.w, 2), 0);
                const int p3 = __dp2a_lo((int)ps.w, (int)byte_pair(sm.z, sm.w, 3), 0);
                if (EXACT) {
                    int4 *dstw = reinterpret_cast<int4 *>(scratch + n * SCR_WORDS + (r * 8 + b) * 20 + r * 8);
                    dstw[0] = make_int4(sum16[n][0] >> 4, sum16[n][1] >> 4, sum16[n][2] >> 4, sum16[n][3] >> 4);
                    dstw[1] = make_int4(sum16[n][4 % (EXACT ? 8 : 1)] >> 4, sum16[n][5 % (EXACT ? 8 : 1)] >> 4,
                                        sum16[n][6 % (EXACT ? 8 : 1)] >> 4, sum16[n][7 % (EXACT ? 8 : 1)] >> 4);
                    dstw[2] = make_int4(p0, p1, p2, p3);
                    dstw[3] = make_int4(__float_as_int(dall), __float_as_int(dmn), 0, 0);
                } else {
                    acc[n] = __fmaf_rn(dall, (float)(sum16[n][0] >> 4), acc[n]);
                    accm[n] = __fmaf_rn(dmn, (float)(p0 + p1 + p2 + p3), accm[n]);
                }
            }
        }
        if (EXACT) {
            __syncwarp();
            // chain phase: lane (r, t = b) walks the superblocks of row r in order, as the AVX2 loop does
#pragma unroll
            for (int n = 0; n < N; ++n) {
                const int *row_scr = scratch + n * SCR_WORDS + r * 8 * 20 + r * 8;
                for (int bb = 0; bb < nsb; ++bb) {
                    const int *e = row_scr + bb * 20;
                    acc[n] = __fmaf_rn(__int_as_float(e[12]), (float)e[b], acc[n]);
                    accm[n] = __fmaf_rn(__int_as_float(e[13]), (float)e[8 + (b & 3)], accm[n]);
                }
            }
        }
        if (ks == ksteps - 1) {                              // row group finished: horizontal sums + store
            const int64_t row = (int64_t)mt.x * TILE_ROWS + r;
#pragma unroll
            for (int n = 0; n < N; ++n) {
                float a = acc[n], m = accm[n];
                if (EXACT) {
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 4));    // hsum_float_8: t and t+4
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 2));    //   (r0+r2), (r1+r3)
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 1));
                    m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 2));    // acc_m + movehl
                    m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 1));    // + movehdup
                    a = __fadd_rn(a, m);
                } else {
                    a = __fadd_rn(a, m);
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 1));
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 2));
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 4));
                }
                if (b == 0 && row < p.M) p.dst[(int64_t)n * p.dst_stride + row] = a;
                acc[n] = 0.0f; accm[n] = 0.0f;
            }
        }
        __syncwarp();
        produce(s);
    }
}

}  // namespace

namespace {

constexpr size_t SMEM_MAX = 232448;      // 227 KB per CTA on sm_100

size_t fixed_smem(int N, int nb, bool exact) {
    size_t off = (size_t)N * nb * Q8_STRIDE;
    off += ((size_t)N * nb * 4 + 127) & ~(size_t)127;
    off = (off + 127) & ~(size_t)127;
    off += 16;                                               // row-group counter
    if (exact) off += (size_t)WARPS * N * SCR_WORDS * 4;
    return off + 128;
}
constexpr size_t PER_STAGE = (size_t)WARPS * (STAGE_BYTES + 16);

template <int N, bool EXACT>
cudaError_t launch_tma_n(TmaGemvParams p, int want_stages, cudaStream_t stream) {
    auto kern = gemv_q4k_tma_kernel<N, EXACT>;
    const size_t fixed = fixed_smem(N, p.nb, EXACT);
    if (fixed + 2 * PER_STAGE > SMEM_MAX) return cudaErrorNotSupported;      // activations do not fit: caller tiles
    int stages = want_stages;
    if (stages <= 0) {
        // two CTAs per SM with 2-3 stages when the activations are small, else one CTA with up to 4
        const size_t half = SMEM_MAX / 2 - 1024;
        if (fixed + 2 * PER_STAGE <= half) stages = (fixed + 3 * PER_STAGE <= half) ? 3 : 2;
        else stages = (int)((SMEM_MAX - fixed) / PER_STAGE) > 4 ? 4 : (int)((SMEM_MAX - fixed) / PER_STAGE);
    }
    while (stages > 2 && fixed + stages * PER_STAGE > SMEM_MAX) --stages;
    p.stages = stages;
    const size_t smem = fixed + stages * PER_STAGE;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    int per_sm = 1;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS * 32, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorNotSupported;
    int64_t grid = (int64_t)sm_count() * per_sm;
    if (grid > p.n_groups) grid = p.n_groups;
    kern<<<(unsigned)grid, WARPS * 32, smem, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}

template <bool EXACT>
cudaError_t launch_tma_e(int N, const TmaGemvParams &p, int stages, cudaStream_t st) {
    switch (N) {
        case 1: return launch_tma_n<1, EXACT>(p, stages, st);
        case 2: return launch_tma_n<2, EXACT>(p, stages, st);
        case 3: return launch_tma_n<3, EXACT>(p, stages, st);
        case 4: return launch_tma_n<4, EXACT>(p, stages, st);
        case 5: return launch_tma_n<5, EXACT>(p, stages, st);
        case 6: return launch_tma_n<6, EXACT>(p, stages, st);
        case 7: return launch_tma_n<7, EXACT>(p, stages, st);
        case 8: return launch_tma_n<8, EXACT>(p, stages, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace

cudaError_t launch_gemv_q4k_q8k_tma(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                    const void *xq, int N, int64_t xq_row_stride_bytes,
                                    float *dst, int64_t dst_col_stride_bytes, bool exact_order, int stages,
                                    cudaStream_t stream) {
    if (M == 0 || N == 0) return cudaSuccess;
    TmaGemvParams p;
    p.w = (const uint8_t *)w; p.M = M; p.w_stride = w_row_stride_bytes; p.nb = (int)(K / QK_K);
    p.xq = (const uint8_t *)xq; p.xq_stride = xq_row_stride_bytes;
    p.dst = dst; p.dst_stride = dst_col_stride_bytes / 4;
    p.n_groups = (M + TILE_ROWS - 1) / TILE_ROWS;
    p.stages = 0;
    return exact_order ? launch_tma_e<true>(N, p, stages, stream) : launch_tma_e<false>(N, p, stages, stream);
}

}  // namespace q4k
