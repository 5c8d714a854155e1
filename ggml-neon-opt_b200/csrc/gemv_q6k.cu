// gemv_q6k.cu — decode GEMV (batch 1..8) for Q6_K weights.  dst[n][m] = vec_dot(W row m, Q8_K(x row n)).
// replaces ggml_vec_dot_q6_K_q8_K (L/ggml/src/ggml-cpu/arch/x86/quants.c:2129-2218, generic
// L/ggml/src/ggml-cpu/quants.c:705-755) under ggml_compute_forward_mul_mat (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401,
// type_traits_cpu[GGML_TYPE_Q6_K], :296-301) for N <= 8, with quantize_row_q8_K (L/ggml/src/ggml-quants.c:2555-2592)
// inside the same launch when the activations arrive as F32.  TinyLlama Q4_K_M keeps attn_v / ffn_down of 10 layers and
// output.weight in Q6_K (L/src/llama-quant.cpp:228-230, 302-303, 358-364).
//
// block_q6_K (L/ggml/src/ggml-common.h:326-336) is 210 bytes: ql[128] @0, qh[64] @128, int8 scales[16] @192, fp16 d @208.
// 210 = 2 mod 4, so every other block of a row starts on a 2-byte boundary and a row of K/256 blocks is 16-byte
// aligned only when K/256 is a multiple of 8 (TinyLlama's ffn_down, K = 5632, is not).
//
// HBM-bound; no tensor cores.  Structure (the bulk-copy ring of gemv_q4k_tma.cu, re-cut for the 210-byte blocks):
//   * each warp owns a ring of S slots; a slot holds a 4-row x 8-superblock tile.  A row piece (<= 1680 B) is fetched
//     with one cp.async.bulk over the 16-byte-aligned span that contains it; the up-to-14-byte shift is kept per row.
//     (The last < 16 bytes of a matrix that does not end on a 16-byte boundary are copied with plain loads, so nothing
//     past the matrix is ever read.)  Row pitch in the slot = 1696 B = 424 words: with the 52.5-word block pitch the 32
//     lanes of a warp fall on 32 different banks.
//   * one lane per superblock (lane (r, b): row r of the tile, superblock b): 53 word reads (LDS.32 + funnel shift by
//     the block's 0 / 2 byte misalignment) bring ql / qh / scales into registers, the slot goes back to the copy engine,
//     arithmetic runs from registers.
//   * integers: the 6-bit levels are rebuilt four at a time WITHOUT the -32 offset (two LOP3 per word),
//     dp4a.u32.s32 against the Q8_K bytes gives the sixteen sub-block sums, scaled by the int8 scales; the offset is
//     folded out exactly through the Q8_K group sums:  sum (q - 32) y = sum q y - 32 bsums  (what the reference's AVX
//     branch does, arch/x86/quants.c:2237-2239).  The per-superblock integer `sumi` equals the reference's.
//   * float: acc += (d_w * d_y) * (float)sumi per superblock and lane in K order, then the eight lanes of a row
//     (xor 1, 2, 4).  Nothing depends on the grid or the ring depth.  Within 1e-5 of ggml_vec_dot_q6_K_q8_K.
//   * Q8_K activations in shared memory once per CTA: 304 B per block (256 qs, 16 bsums, 16 pad): 76 words = 12 banks
//     per block step, conflict-free for the eight blocks of a tile step; the four rows broadcast.
//   * 4-row groups are dealt round-robin to the CTAs (the grid sweeps the matrix front to back), warps of a CTA pull
//     from a shared counter; programmatic dependent launch: weights start moving before griddepcontrol.wait.
#include <cstdlib>

#include "q4k_common.cuh"
#include "q4k_kernels.h"

#include "gemv_stage.cuh"

namespace q4k {

namespace {

constexpr int BLK_Q6K = 210;
constexpr int Q6_TILE_ROWS = 4;
constexpr int Q6_TILE_SB = 8;
constexpr int Q6_PIECE = Q6_TILE_SB * BLK_Q6K;         // 1680: a row piece, a multiple of 16
constexpr int Q6_ROWP = 1696;                          // slot row pitch: piece + worst shift (14) rounded up to 16
constexpr int Q6_SLOT = Q6_TILE_ROWS * Q6_ROWP;        // 6784
constexpr int Q6_Q8P = 304;                            // Q8_K block in shared memory: 256 qs + 16 x int16 bsums + pad
constexpr int Q6_WARPS = 8;

struct Q6Params {
    const uint8_t *w;
    int64_t M, w_stride;
    int nb;
    const uint8_t *xq;      // Q8_K activations, or
    const float *xf;        // F32 activations (quantized in the kernel) when xq == nullptr
    int64_t x_stride;
    float *dst;
    int64_t dst_stride;     // floats between output columns
    int64_t n_groups;
    int stages;
    int32_t *dump;          // DUMP instantiation: sumi[(n * M + row) * nb + block]
};

__device__ __forceinline__ int sx8(uint32_t w, int i) { return (int)(int8_t)(w >> (8 * i)); }
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ unsigned short lds16(uint32_t a) {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return v;
}

template <int N, bool DUMP>
__global__ void __launch_bounds__(Q6_WARPS * 32, 2)
gemv_q6k_kernel(const __grid_constant__ Q6Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int nb = p.nb, S = p.stages;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r = lane >> 3, b = lane & 7;

    size_t off = 0;
    uint8_t *q8_s = smem;                                  off += (size_t)N * nb * Q6_Q8P;
    float *d_s = reinterpret_cast<float *>(smem + off);    off += ((size_t)N * nb * 4 + 127) & ~(size_t)127;
    uint8_t *wbuf = smem + off + (size_t)warp * S * Q6_SLOT;       off += (size_t)Q6_WARPS * S * Q6_SLOT;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + off) + warp * S;   off += (size_t)Q6_WARPS * S * 8;
    int2 *meta = reinterpret_cast<int2 *>(smem + off) + warp * S;  off += (size_t)Q6_WARPS * S * 8;
    int *counter = reinterpret_cast<int *>(smem + off);

    const int ksteps = (nb + Q6_TILE_SB - 1) / Q6_TILE_SB;
    const uint8_t *w_end_ok = reinterpret_cast<const uint8_t *>(
        reinterpret_cast<uintptr_t>(p.w + (p.M - 1) * p.w_stride + (int64_t)nb * BLK_Q6K) & ~(uintptr_t)15);

    if (threadIdx.x == 0) *counter = 0;
    if (lane == 0) {
        for (int s = 0; s < S; ++s) mbar_init(smem_u32(&bars[s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // ---- producer state (warp-uniform) ----
    int64_t cur_rg = -1;
    int prod_ks = ksteps;
    bool exhausted = false;
    auto produce = [&](int s) {
        if (!exhausted && prod_ks == ksteps) {
            int nxt = 0;
            if (lane == 0) nxt = atomicAdd(counter, 1);
            nxt = __shfl_sync(0xffffffffu, nxt, 0);
            cur_rg = (int64_t)blockIdx.x + (int64_t)nxt * gridDim.x;
            prod_ks = 0;
            if (cur_rg >= p.n_groups) exhausted = true;
        }
        const uint32_t bar = smem_u32(&bars[s]);
        if (exhausted) {
            if (lane == 0) {
                meta[s] = make_int2(-1, 0);
                mbar_arrive(bar);
            }
            return;
        }
        const int nsb = min(Q6_TILE_SB, nb - prod_ks * Q6_TILE_SB);
        // lanes 0..3: the aligned span of their row piece
        int64_t row = cur_rg * Q6_TILE_ROWS + (lane & 3);
        if (row >= p.M) row = p.M - 1;                     // ragged last group: stay in bounds, result is not stored
        const uint8_t *src = p.w + row * p.w_stride + (int64_t)prod_ks * Q6_PIECE;
        const uint8_t *a = reinterpret_cast<const uint8_t *>(reinterpret_cast<uintptr_t>(src) & ~(uintptr_t)15);
        const uint32_t want = (uint32_t)(src - a) + (uint32_t)nsb * BLK_Q6K;       // bytes of the span that matter
        uint32_t len = (want + 15u) & ~15u;
        if (a + len > w_end_ok) len = w_end_ok > a ? (uint32_t)(w_end_ok - a) : 0u;  // never read past the matrix
        uint32_t total = lane < Q6_TILE_ROWS ? len : 0u;
        total += __shfl_xor_sync(0xffffffffu, total, 1);
        total += __shfl_xor_sync(0xffffffffu, total, 2);
        if (lane == 0) {
            meta[s] = make_int2((int)cur_rg, prod_ks);
            mbar_expect_tx(bar, total);
        }
        __syncwarp();
        if (lane < Q6_TILE_ROWS) {
            uint8_t *dstp = wbuf + (size_t)s * Q6_SLOT + lane * Q6_ROWP;
            for (uint32_t i = len; i < want; ++i) dstp[i] = a[i];                  // the matrix tail (< 16 bytes), plain loads
            fence_proxy_async();                      // the warp's generic accesses of this slot are done (after __syncwarp)
            if (len) bulk_g2s(smem_u32(dstp), a, len, bar);
        }
        __syncwarp();                                 // the tail bytes stored by lanes 0..3 are visible to the lanes that read them
        ++prod_ks;
    };

    for (int s = 0; s < S; ++s) produce(s);
    griddep_launch_dependents();
    griddep_wait();
    // ---- activations -> shared memory ----
    if (p.xq) {
        for (int i = threadIdx.x; i < N * nb * 72; i += blockDim.x) {
            const int e = i / 72, wd = i - e * 72, n = e / nb, kb = e - n * nb;
            const uint8_t *src = p.xq + n * p.x_stride + (int64_t)kb * BLK_Q8K;
            reinterpret_cast<uint32_t *>(q8_s + (size_t)e * Q6_Q8P)[wd] = reinterpret_cast<const uint32_t *>(src + Q8K_QS)[wd];
        }
        for (int i = threadIdx.x; i < N * nb; i += blockDim.x) {
            const int n = i / nb, kb = i - n * nb;
            d_s[i] = *reinterpret_cast<const float *>(p.xq + n * p.x_stride + (int64_t)kb * BLK_Q8K);
        }
    } else {
        // three blocks per warp in flight: all loads of a round trip before any arithmetic (the quantizer is ~1000 dependent cycles)
        for (int e0 = warp; e0 < N * nb; e0 += 3 * Q6_WARPS) {
            float4 f0[3], f1[3];
#pragma unroll
            for (int rr = 0; rr < 3; ++rr) {
                const int e = e0 + rr * Q6_WARPS;
                if (e < N * nb) {                                            // warp-uniform
                    const int n = e / nb, kb = e - n * nb;
                    const float4 *src = reinterpret_cast<const float4 *>(reinterpret_cast<const char *>(p.xf) + n * p.x_stride) + kb * 64;
                    f0[rr] = src[lane];
                    f1[rr] = src[32 + lane];
                }
            }
#pragma unroll
            for (int rr = 0; rr < 3; ++rr) {
                const int e = e0 + rr * Q6_WARPS;
                if (e < N * nb) {
                    const float v[8] = {f0[rr].x, f0[rr].y, f0[rr].z, f0[rr].w, f1[rr].x, f1[rr].y, f1[rr].z, f1[rr].w};
                    uint32_t w_lo, w_hi;
                    int s_lo, s_hi;
                    const float d = quantize_block_q8_K<4>(v, lane, w_lo, w_hi, s_lo, s_hi);
                    uint8_t *dq = q8_s + (size_t)e * Q6_Q8P;
                    reinterpret_cast<uint32_t *>(dq)[lane] = w_lo;
                    reinterpret_cast<uint32_t *>(dq)[32 + lane] = w_hi;
                    if ((lane & 3) == 0) {
                        reinterpret_cast<int16_t *>(dq + 256)[lane >> 2] = (int16_t)s_lo;
                        reinterpret_cast<int16_t *>(dq + 256)[8 + (lane >> 2)] = (int16_t)s_hi;
                    }
                    if (lane == 0) d_s[e] = d;
                }
            }
        }
    }
    __syncthreads();

    float acc[N];
#pragma unroll
    for (int n = 0; n < N; ++n) acc[n] = 0.0f;

    int s = 0;
    uint32_t parity = 0;
    for (;;) {
        mbar_wait(smem_u32(&bars[s]), parity);
        const int2 mt = meta[s];
        if (mt.x < 0) break;
        const int ks = mt.y;
        const int nsb = min(Q6_TILE_SB, nb - ks * Q6_TILE_SB);
        const bool active = b < nsb;
        const int64_t row = (int64_t)mt.x * Q6_TILE_ROWS + r;
        const int64_t rowc = row < p.M ? row : p.M - 1;

        // the lane's superblock: shared memory -> registers (word j = bytes 4j..4j+3 of the block)
        uint32_t ql[32], qh[16], scw[4];
        float dw;
        {
            const uint32_t shift = (uint32_t)(reinterpret_cast<uintptr_t>(p.w + rowc * p.w_stride + (int64_t)ks * Q6_PIECE) & 15u);
            const uint8_t *blk = wbuf + (size_t)s * Q6_SLOT + r * Q6_ROWP + shift + (active ? b : 0) * BLK_Q6K;
            const uint32_t sa = smem_u32(blk), al = sa & ~3u;          // explicit ld.shared: the address arithmetic hides the state space
            const uint32_t fs = (sa & 3u) * 8u;                        // 0 or 16
            uint32_t prev = lds32(al);
#pragma unroll
            for (int j = 0; j < 52; ++j) {
                const uint32_t next = lds32(al + 4 * (j + 1));
                const uint32_t v = __funnelshift_r(prev, next, fs);
                if (j < 32) ql[j] = v;
                else if (j < 48) qh[j - 32] = v;
                else scw[j - 48] = v;
                prev = next;
            }
            dw = __half2float(__ushort_as_half(lds16(sa + 208)));
        }
        __syncwarp();
        produce(s);

        if (active) {
            const int kb = ks * Q6_TILE_SB + b;
            int sumi[N];
#pragma unroll
            for (int n = 0; n < N; ++n) sumi[n] = 0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
#pragma unroll
                for (int wq = 0; wq < 2; ++wq) {
                    uint32_t v[4][4];                  // [t][i]: levels (0..63) of elements 128h + 32t + 16wq + 4i ..+3
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const uint32_t a = ql[16 * h + 4 * wq + i], c = ql[16 * h + 8 + 4 * wq + i], hh = qh[8 * h + 4 * wq + i];
                        v[0][i] = (a & 0x0f0f0f0fu) | ((hh << 4) & 0x30303030u);
                        v[1][i] = (c & 0x0f0f0f0fu) | ((hh << 2) & 0x30303030u);
                        v[2][i] = ((a >> 4) & 0x0f0f0f0fu) | (hh & 0x30303030u);
                        v[3][i] = ((c >> 4) & 0x0f0f0f0fu) | ((hh >> 2) & 0x30303030u);
                    }
#pragma unroll
                    for (int n = 0; n < N; ++n) {
                        const uint8_t *yb = q8_s + (size_t)(n * nb + kb) * Q6_Q8P + 128 * h + 16 * wq;
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const uint4 y = lds128(yb + 32 * t);
                            int pj = dp4a_u8s8(v[t][0], y.x, 0);
                            pj = dp4a_u8s8(v[t][1], y.y, pj);
                            pj = dp4a_u8s8(v[t][2], y.z, pj);
                            pj = dp4a_u8s8(v[t][3], y.w, pj);
                            const int j = 8 * h + 2 * t + wq;           // sub-block of 16
                            sumi[n] += sx8(scw[j >> 2], j & 3) * pj;
                        }
                    }
                }
            }
#pragma unroll
            for (int n = 0; n < N; ++n) {
                const uint8_t *yb = q8_s + (size_t)(n * nb + kb) * Q6_Q8P + 256;
                const uint4 b0 = lds128(yb), b1 = lds128(yb + 16);
                const uint32_t bs[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                int corr = 0;                                            // sum_j scales[j] * bsums[j]
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    corr = (k & 1) ? __dp2a_hi((int)bs[k], (int)scw[k >> 1], corr) : __dp2a_lo((int)bs[k], (int)scw[k >> 1], corr);
                const int total = sumi[n] - 32 * corr;
                if (DUMP) { if (row < p.M) p.dump[((int64_t)n * p.M + row) * nb + kb] = total; }
                acc[n] = __fmaf_rn(__fmul_rn(d_s[n * nb + kb], dw), (float)total, acc[n]);
            }
        }
        if (ks == ksteps - 1) {
#pragma unroll
            for (int n = 0; n < N; ++n) {
                float a = acc[n];
                a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 1));
                a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 2));
                a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 4));
                if (b == 0 && row < p.M) p.dst[(int64_t)n * p.dst_stride + row] = a;
                acc[n] = 0.0f;
            }
        }
        if (++s == S) { s = 0; parity ^= 1u; }
    }
}

constexpr size_t SMEM_MAX_Q6 = 232448;
constexpr size_t PER_STAGE_Q6 = (size_t)Q6_WARPS * (Q6_SLOT + 16);

size_t fixed_smem_q6(int N, int nb) {
    return (size_t)N * nb * Q6_Q8P + (((size_t)N * nb * 4 + 127) & ~(size_t)127) + 16 + 128;
}

template <int N, bool DUMP>
cudaError_t launch_q6_n(Q6Params p, bool pdl, cudaStream_t stream) {
    auto kern = gemv_q6k_kernel<N, DUMP>;
    const size_t fixed = fixed_smem_q6(N, p.nb);
    if (fixed + PER_STAGE_Q6 > SMEM_MAX_Q6) return cudaErrorNotSupported;        // activations do not fit: caller tiles
    int stages = 2;
    if (knobs().q6k_stages > 0) stages = knobs().q6k_stages;
    while (stages > 1 && 2 * (fixed + stages * PER_STAGE_Q6 + 1024) > SMEM_MAX_Q6) --stages;
    if (fixed + stages * PER_STAGE_Q6 > SMEM_MAX_Q6) stages = (int)((SMEM_MAX_Q6 - fixed) / PER_STAGE_Q6);
    p.stages = stages;
    const size_t smem = fixed + stages * PER_STAGE_Q6;
    if (const cudaError_t ea = ensure_max_dynamic_smem((const void *)kern, (int)SMEM_MAX_Q6)) return ea;
    const int per_sm = (2 * (smem + 1024) <= SMEM_MAX_Q6) ? 2 : 1;
    int64_t grid = (int64_t)sm_count() * per_sm;
    if (grid > p.n_groups) grid = p.n_groups;
    if (grid < 1) grid = 1;

    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(Q6_WARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
    count_launch();
    return e;
}

template <bool DUMP>
cudaError_t launch_q6_d(int N, const Q6Params &p, bool pdl, cudaStream_t st) {
    switch (N) {
        case 1: return launch_q6_n<1, DUMP>(p, pdl, st);
        case 2: return launch_q6_n<2, DUMP>(p, pdl, st);
        case 3: return launch_q6_n<3, DUMP>(p, pdl, st);
        case 4: return launch_q6_n<4, DUMP>(p, pdl, st);
        case 5: return launch_q6_n<5, DUMP>(p, pdl, st);
        case 6: return launch_q6_n<6, DUMP>(p, pdl, st);
        case 7: return launch_q6_n<7, DUMP>(p, pdl, st);
        case 8: return launch_q6_n<8, DUMP>(p, pdl, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace

// xq != nullptr: activations are Q8_K blocks; else xf: F32 rows, quantized inside the kernel.  dump != nullptr: also
// write the per-superblock integer sums (parity tooling; same integer code as the product path).
// cudaErrorNotSupported when N columns of K activations do not fit in shared memory (the caller splits the columns).
cudaError_t launch_gemv_q6k(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                            const void *xq, const float *xf, int N, int64_t x_row_stride_bytes,
                            float *dst, int64_t dst_col_stride_bytes, int32_t *dump, bool pdl, cudaStream_t stream) {
    if (M == 0 || N == 0) return cudaSuccess;
    Q6Params p;
    p.w = (const uint8_t *)w; p.M = M; p.w_stride = w_row_stride_bytes; p.nb = (int)(K / QK_K);
    p.xq = (const uint8_t *)xq; p.xf = xf; p.x_stride = x_row_stride_bytes;
    p.dst = dst; p.dst_stride = dst_col_stride_bytes / 4;
    p.n_groups = (M + Q6_TILE_ROWS - 1) / Q6_TILE_ROWS;
    if (p.n_groups > 0x7fffffff) return cudaErrorInvalidValue;
    p.stages = 0; p.dump = dump;
    return dump ? launch_q6_d<true>(N, p, pdl, stream) : launch_q6_d<false>(N, p, pdl, stream);
}

}  // namespace q4k
