// gemv_q4k_wide.cu — batched decode (up to 32 columns per launch; wider batches in column groups of 32), integer dot
// products on the tensor pipe.  Replaces the vec_dot loop of ggml_compute_forward_mul_mat_one_chunk
// (L/ggml/src/ggml-cpu/ggml-cpu.c:1115-1203) + ggml_vec_dot_q4_K_q8_K (L/ggml/src/ggml-cpu/arch/x86/quants.c:1741-1823) for
// the batch sizes between the decode kernel (gemv_q4k_mma.cu, N <= 8: every CTA holds ALL activations in shared memory) and
// the tcgen05 GEMM (gemm_q4k_tc.cu: a 128-column tile whose time does not shrink with N — 4096 x 14336 x 32 took 41 us against
// ~5 us of HBM time).  At these N the op is still HBM-bound (~106 int8 op per weight byte at N = 32): the weights must stream
// ONCE, and what has to be solved is where the N x K activations live — 32 columns of K = 14336 are 523 KB of Q8_K, more than
// two SMs' shared memory.
//
// Design.  The activations are quantized once per call (quantize_q8_K_wide_kernel: the arithmetic of quantize_q8k.cu, written
// in the kernel's own staged layout) and then STREAM through shared memory next to the weights:
//   * a round is 4 superblocks of K for TS 16-row tiles (a "pass" = those TS tiles over all of K).  Per round ONE lane issues
//     TS + 1 copies on one mbarrier: TS TMA boxes (16 rows x 608 B of raw block_q4_K through a 2-D tensor map over the weight
//     matrix — the 32 bytes beyond the round's 576 are the row padding the fragment loads need; rows past M and bytes past
//     the row read as zero) and one bulk copy of the round's activation stage, which the quantizer has laid out in global
//     memory as the exact shared-memory image.  (Measured: an SM retires one bulk copy per ~80 cycles whatever its size — 48
//     copies of 576-1152 B per round held the first version of this kernel to 2.1 TB/s of weights alone.)
//   * sixteen consumer warps = TS tiles x 4 superblocks x CH column halves (TS x CH = 4).  A warp-step is mma.sync.m16n8k32
//     (u8 x s8 -> s32) over the warp's NTW column tiles of 8 with the A fragments (weights) loaded once: the "columns are batch
//     columns" form of gemv_q4k_mma.cu, C[row][column] = the sub-block dot product p[j].  Sixteen warps, not eight: a
//     warp-step is a chain of shared-memory loads, MMAs and integer multiply-adds, and two warps per scheduler left the
//     arithmetic un-overlapped with the copies (28672 x 8192 N = 16: 27 of 56 us; profiles/r2_wide_sweep_v1_8warps.txt).
//   * at the end of a pass the warps' partial sums meet in shared memory, are added in superblock-class order and stored.
//     Activations are re-streamed once per pass (from L2: the staged copy of 32 columns x K = 14336 is 530 KB), weights
//     cross HBM once.
// Shapes: n_pad = 8: TS = 4, CH = 1 (TS = 2 and eight warps when the matrix has fewer than 4 x 148 tiles); n_pad = 16: TS = 2, CH = 2, one column tile per warp; n_pad = 24 / 32: TS = 2, CH = 2, two
// column tiles per warp (17..24 columns are padded to 32 with zero columns).
// Staged activation block (288 B, 16-byte aligned): 256 int8 | 8 x int16 pair sums (bsums[2j] + bsums[2j+1]) | float d | pad.
// Shared-memory pitches: activation column 4 x 288 + 32 B (8 words mod 32) and weight row 576 + 32 B (24 words mod 32): the
// 8-byte fragment loads of a half-warp (4 rows or columns x 4 pieces) hit 16 distinct bank pairs.
//
// Integers are exact (same sumi / summ as the reference: DUMP instantiation, q4k_mul_mat_dump).  Float association: per
// (row, column) every warp adds its superblocks (kb = s, s + 4, ...) in K order as  acc += (d * dy) * sumi,
// accm += (-dmin * dy) * summ;  the result is the four superblock classes' (acc + accm) added in class order.  Independent of
// grid, ring depth, launch shape and of N.  Within 1e-5 of ggml_vec_dot_q4_K_q8_K (tests/test_gpu_wide.py).
#include <algorithm>

#include <cuda.h>

#include "gemv_stage.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

constexpr int WROWS = 16;                       // rows per tile (the MMA's m)
constexpr int RSB = 4;                          // superblocks per round
// consumer warps = TS x RSB x CH (16, or 8 for the small-matrix shape of n_pad = 8); + the producer warp
constexpr int XBLK = 288;                       // staged activation block
constexpr int XD_OFF = 272;                     // float d inside it (pair sums at 256)
constexpr int APITCH = RSB * XBLK + 32;         // activation column pitch in a stage (1184 B = 296 words = 8 mod 32)
constexpr int WPITCH = RSB * BLK_Q4K + 32;      // weight row pitch in a slot (608 B = 152 words = 24 mod 32)
constexpr int WSLOT = WROWS * WPITCH;           // 9728 B
constexpr size_t WIDE_SMEM_MAX = 232448;

struct WideParams {
    CUtensorMap wmap;           // the weights as a 2-D tensor of 8-byte elements [M][row bytes / 8], box = 16 rows x 608 B
    int64_t M;
    int nb;
    const uint8_t *xs;          // staged activations: [rounds][n_pad] column pieces of APITCH bytes (4 blocks of XBLK + pad) = the stage images
    float *dst;
    int64_t dst_stride;         // floats between result columns
    int N, n_pad;
    int64_t tiles, passes;      // 16-row tiles, groups of TS tiles
    int stages;                 // ring depth
    int dbg;                    // development (knob wide_dbg): 1 = no activation copies, 2 = no arithmetic, 4 = no weight copies — wrong results, timing only
    int32_t *dump_p, *dump_sumi, *dump_summ;
};

__device__ __forceinline__ void imma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint2 lds64w(const uint8_t *p) { return *reinterpret_cast<const uint2 *>(p); }
__device__ __forceinline__ void bar_sync_named(int id, int threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory"); }

// summ of one (row, column): sum_j m[j] * ps[j]; mz / mw = the row's eight mins as bytes, ps = the column's pair sums
__device__ __forceinline__ int mins_dot(const uint4 ps, uint32_t mz, uint32_t mw) {
    return __dp2a_hi((int)ps.w, (int)mw, __dp2a_lo((int)ps.z, (int)mw, __dp2a_hi((int)ps.y, (int)mz, __dp2a_lo((int)ps.x, (int)mz, 0))));
}

// TS tiles per pass, CH column halves, NTW column tiles of 8 per warp; RAG: the last column half has fewer than NTW tiles
template <int TS, int CH, int NTW, bool RAG, bool DUMP>
__global__ void __launch_bounds__((TS * CH * RSB + 1) * 32, 1)
gemv_q4k_wide_kernel(const __grid_constant__ WideParams p) {
    constexpr int CW = TS * CH * RSB;
    static_assert(CW == 16 || CW == 8, "eight or sixteen consumer warps");
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nb = p.nb, S = p.stages;
    const int rounds = (nb + RSB - 1) / RSB;
    const uint32_t abytes = (uint32_t)p.n_pad * APITCH;                      // activation part of a stage
    const size_t stage = (size_t)abytes + (size_t)TS * WSLOT;               // + TS weight boxes

    uint8_t *ring = smem;
    float *red_s = reinterpret_cast<float *>(smem + (size_t)S * stage);      // [warp][NTW * 4][lane]
    uint64_t *full = reinterpret_cast<uint64_t *>(red_s + CW * NTW * 128);
    uint64_t *empty = full + S;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full[s]), 1); mbar_init(smem_u32(&empty[s]), CW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    griddep_launch_dependents();

    // passes of this CTA: tile groups blockIdx.x, blockIdx.x + gridDim.x, ... (the grid sweeps the matrix front to back)
    const int64_t my_passes = p.passes > blockIdx.x ? (p.passes - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total = my_passes * rounds;                  // rounds this CTA walks through

    if (warp == CW) {
        // ---- producer: one lane, TS + 1 copies per round on the stage's barrier ----
        if (lane != 0) return;
        int64_t pass = blockIdx.x;
        int rd = 0, slot = 0;
        uint32_t ph = 0u;
        bool waited = false;                                   // activations depend on the quantizer in front of this kernel
        for (int64_t q = 0; q < total; ++q) {
            mbar_wait(smem_u32(&empty[slot]), ph ^ 1u);        // (falls through for the first S rounds)
            const uint32_t bar = smem_u32(&full[slot]);
            uint8_t *st = ring + (size_t)slot * stage;
            const uint32_t tx = ((p.dbg & 1) ? 0u : abytes) + ((p.dbg & 4) ? 0u : (uint32_t)(TS * WSLOT));
            if (tx) mbar_expect_tx(bar, tx); else mbar_arrive(bar);
            fence_proxy_async();
            if (!(p.dbg & 4)) {
#pragma unroll
                for (int ts = 0; ts < TS; ++ts)
                    tma_box_g2s(smem_u32(st + abytes + (size_t)ts * WSLOT), &p.wmap, rd * (RSB * BLK_Q4K / 8), (int)((pass * TS + ts) * WROWS), bar);
            }
            if (!waited) { griddep_wait(); waited = true; }    // the weights of the first round are already on their way
            if (!(p.dbg & 1)) bulk_g2s(smem_u32(st), p.xs + (size_t)rd * abytes, abytes, bar);
            if (++rd == rounds) { rd = 0; pass += gridDim.x; }
            if (++slot == S) { slot = 0; ph ^= 1u; }
        }
        return;
    }

    // ---- consumer warps: warp -> (tile slot, column half, superblock of the round) ----
    const int sb = warp & (RSB - 1), ch = (warp >> 2) % CH, ts = warp / (RSB * CH);
    const int r = lane >> 2, c = lane & 3;
    const int t0 = ch * NTW;                                    // first column tile of this warp
    const int nt_mine = RAG ? min(NTW, p.n_pad / 8 - t0) : NTW; // column tiles that exist (n_pad = 24: the second half has one)
    int slot = 0;
    uint32_t ph = 0u;
    for (int64_t ip = 0; ip < my_passes; ++ip) {
        const int64_t pass = blockIdx.x + ip * gridDim.x;
        const int64_t tile = pass * TS + ts;
        const bool tile_ok = tile < p.tiles;
        float acc[NTW][4], accm[NTW][4];
#pragma unroll
        for (int t = 0; t < NTW; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) { acc[t][i] = 0.0f; accm[t][i] = 0.0f; }

        for (int rd = 0; rd < rounds; ++rd) {
            mbar_wait(smem_u32(&full[slot]), ph);
            const int kb = rd * RSB + sb;
            if (kb < nb && tile_ok && !(p.dbg & 2)) {
                const uint8_t *st = ring + (size_t)slot * stage;
                const uint8_t *p0 = st + abytes + (size_t)ts * WSLOT + r * WPITCH + sb * BLK_Q4K, *p1 = p0 + 8 * WPITCH;
                const uint8_t *ya = st + (size_t)(8 * t0) * APITCH + sb * XBLK;
                const uint4 h0 = lds128(p0), h1 = lds128(p1);
                const uint4 u0 = unpack_scales_mins(h0), u1 = unpack_scales_mins(h1);
                int s[NTW][4];
#pragma unroll
                for (int t = 0; t < NTW; ++t) { s[t][0] = 0; s[t][1] = 0; s[t][2] = 0; s[t][3] = 0; }
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const uint2 wq = lds64w(p0 + 16 + 32 * g + 8 * c), vq = lds64w(p1 + 16 + 32 * g + 8 * c);
                    const uint32_t a0 = wq.x & 0x0f0f0f0fu, a1 = vq.x & 0x0f0f0f0fu, a2 = wq.y & 0x0f0f0f0fu, a3 = vq.y & 0x0f0f0f0fu;
                    const uint32_t e0 = wq.x & 0xf0f0f0f0u, e1 = vq.x & 0xf0f0f0f0u, e2 = wq.y & 0xf0f0f0f0u, e3 = vq.y & 0xf0f0f0f0u;
                    const uint32_t q0 = (g < 2 ? u0.x : u0.y) >> (16 * (g & 1)), q1 = (g < 2 ? u1.x : u1.y) >> (16 * (g & 1));
                    const int lo0 = (int)((q0 & 0xffu) << 4), hi0 = (int)((q0 >> 8) & 0xffu);
                    const int lo1 = (int)((q1 & 0xffu) << 4), hi1 = (int)((q1 >> 8) & 0xffu);
#pragma unroll
                    for (int t = 0; t < NTW; ++t) {
                        if (!RAG || t < nt_mine) {                                                    // warp-uniform
                            const uint8_t *ycol = ya + (size_t)(8 * t + r) * APITCH + 64 * g + 8 * c;     // this lane's B column = lane / 4
                            const uint2 bl = lds64w(ycol), bh = lds64w(ycol + 32);
                            int cl[4] = {0, 0, 0, 0}, cq[4] = {0, 0, 0, 0};
                            imma(cl, a0, a1, a2, a3, bl.x, bl.y);
                            imma(cq, e0, e1, e2, e3, bh.x, bh.y);
                            s[t][0] += lo0 * cl[0] + hi0 * cq[0];
                            s[t][1] += lo0 * cl[1] + hi0 * cq[1];
                            s[t][2] += lo1 * cl[2] + hi1 * cq[2];
                            s[t][3] += lo1 * cl[3] + hi1 * cq[3];
                            if constexpr (DUMP) {      // C[row][column] of this MMA pair IS p[2g] (and 16 x p[2g + 1]) of that (row, column)
#pragma unroll
                                for (int i = 0; i < 4; ++i) {
                                    const int64_t row = tile * WROWS + r + 8 * (i >> 1);
                                    const int col = 8 * (t0 + t) + 2 * c + (i & 1);
                                    if (row < p.M && col < p.N) {
                                        int32_t *o = p.dump_p + (((int64_t)col * p.M + row) * nb + kb) * 8 + 2 * g;
                                        o[0] = cl[i]; o[1] = cq[i] >> 4;
                                    }
                                }
                            }
                        }
                    }
                }
                const float2 dd0 = __half22float2(*reinterpret_cast<const __half2 *>(&h0.x));     // {d, dmin} of row r
                const float2 dd1 = __half22float2(*reinterpret_cast<const __half2 *>(&h1.x));     // row r + 8
#pragma unroll
                for (int t = 0; t < NTW; ++t) {
                    if (!RAG || t < nt_mine) {
                        const uint8_t *ylo = ya + (size_t)(8 * t + 2 * c) * APITCH, *yhi = ylo + APITCH;   // columns 2c, 2c + 1 of the tile
                        const uint4 pl = lds128(ylo + 256), pq = lds128(yhi + 256);
                        const float yd_lo = *reinterpret_cast<const float *>(ylo + XD_OFF), yd_hi = *reinterpret_cast<const float *>(yhi + XD_OFF);
                        const int ml0 = mins_dot(pl, u0.z, u0.w), mh0 = mins_dot(pq, u0.z, u0.w);
                        const int ml1 = mins_dot(pl, u1.z, u1.w), mh1 = mins_dot(pq, u1.z, u1.w);
                        if constexpr (DUMP) {
                            const int si[4] = {s[t][0] >> 4, s[t][1] >> 4, s[t][2] >> 4, s[t][3] >> 4}, sm[4] = {ml0, mh0, ml1, mh1};
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const int64_t row = tile * WROWS + r + 8 * (i >> 1);
                                const int col = 8 * (t0 + t) + 2 * c + (i & 1);
                                if (row < p.M && col < p.N) {
                                    p.dump_sumi[((int64_t)col * p.M + row) * nb + kb] = si[i];
                                    p.dump_summ[((int64_t)col * p.M + row) * nb + kb] = sm[i];
                                }
                            }
                        }
                        acc[t][0] = __fmaf_rn(__fmul_rn(yd_lo, dd0.x), (float)(s[t][0] >> 4), acc[t][0]);
                        acc[t][1] = __fmaf_rn(__fmul_rn(yd_hi, dd0.x), (float)(s[t][1] >> 4), acc[t][1]);
                        acc[t][2] = __fmaf_rn(__fmul_rn(yd_lo, dd1.x), (float)(s[t][2] >> 4), acc[t][2]);
                        acc[t][3] = __fmaf_rn(__fmul_rn(yd_hi, dd1.x), (float)(s[t][3] >> 4), acc[t][3]);
                        accm[t][0] = __fmaf_rn(__fmul_rn(-yd_lo, dd0.y), (float)ml0, accm[t][0]);
                        accm[t][1] = __fmaf_rn(__fmul_rn(-yd_hi, dd0.y), (float)mh0, accm[t][1]);
                        accm[t][2] = __fmaf_rn(__fmul_rn(-yd_lo, dd1.y), (float)ml1, accm[t][2]);
                        accm[t][3] = __fmaf_rn(__fmul_rn(-yd_hi, dd1.y), (float)mh1, accm[t][3]);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&empty[slot]));
            if (++slot == S) { slot = 0; ph ^= 1u; }
        }

        // ---- end of the pass: the four superblock classes of each (tile, column half) meet, in class order ----
        float *mine = red_s + (size_t)warp * NTW * 128;
#pragma unroll
        for (int t = 0; t < NTW; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) mine[(t * 4 + i) * 32 + lane] = __fadd_rn(acc[t][i], accm[t][i]);
        bar_sync_named(1, CW * 32);
        if (sb == 0 && tile_ok) {                                  // one warp per (tile, column half); warp + k = class k's warp
#pragma unroll
            for (int t = 0; t < NTW; ++t) {
                if (!RAG || t < nt_mine) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float *pp = mine + (t * 4 + i) * 32 + lane;
                        float a = pp[0];
#pragma unroll
                        for (int k = 1; k < RSB; ++k) a = __fadd_rn(a, pp[(size_t)k * NTW * 128]);
                        const int64_t row = tile * WROWS + r + 8 * (i >> 1);
                        const int col = 8 * (t0 + t) + 2 * c + (i & 1);
                        if (row < p.M && col < p.N) p.dst[(int64_t)col * p.dst_stride + row] = a;
                    }
                }
            }
        }
        bar_sync_named(1, CW * 32);                                // red_s is free again
    }
}

// Q8_K quantizer into the staged layout — the shared-memory image of every round's activation stage, one after the other:
// [round][column] pieces of APITCH bytes (the round's 4 blocks of the column + 32 B of bank padding).  One warp per
// (column, superblock); columns >= N (padding up to a multiple of 8) are written as zero blocks.
// Same arithmetic as quantize_q8k.cu (quantize_block_q8_K: explicit single-rounding intrinsics).
__global__ void __launch_bounds__(256)
quantize_q8_K_wide_kernel(const float *__restrict__ x, uint8_t *__restrict__ xs, int nb, int N, int n_pad, int64_t x_row_stride_bytes) {
    griddep_launch_dependents();
    griddep_wait();
    const int lane = threadIdx.x & 31;
    const int64_t blk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (blk >= (int64_t)n_pad * nb) return;
    const int64_t col = blk / nb, b = blk - col * nb;
    uint8_t *dst = xs + ((b / RSB) * n_pad + col) * APITCH + (b % RSB) * XBLK;      // round b / 4, column col, block b % 4 of the piece
    uint32_t w_lo = 0u, w_hi = 0u;
    int s_lo = 0, s_hi = 0;
    float d = 0.0f;
    if (col < N) {
        const float *src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + col * x_row_stride_bytes) + b * QK_K;
        const float4 a = ld_act16(src + 4 * lane), e = ld_act16(src + 128 + 4 * lane);
        const float v[8] = {a.x, a.y, a.z, a.w, e.x, e.y, e.z, e.w};
        d = quantize_block_q8_K<8>(v, lane, w_lo, w_hi, s_lo, s_hi);
    }
    reinterpret_cast<uint32_t *>(dst)[lane] = w_lo;
    reinterpret_cast<uint32_t *>(dst)[32 + lane] = w_hi;
    if ((lane & 7) == 0) {
        reinterpret_cast<int16_t *>(dst + 256)[lane >> 3] = (int16_t)s_lo;
        reinterpret_cast<int16_t *>(dst + 256)[4 + (lane >> 3)] = (int16_t)s_hi;
    }
    if (lane == 0) *reinterpret_cast<float *>(dst + XD_OFF) = d;
    if (lane >= 1 && lane <= 3) reinterpret_cast<uint32_t *>(dst + XD_OFF)[lane] = 0u;      // the pad: never read, kept defined
}

template <int TS, int CH, int NTW, bool RAG>
cudaError_t launch_wide_cfg(WideParams p, bool pdl, cudaStream_t stream) {
    constexpr int CW = TS * CH * RSB;
    auto kern = p.dump_sumi ? gemv_q4k_wide_kernel<TS, CH, NTW, RAG, true> : gemv_q4k_wide_kernel<TS, CH, NTW, RAG, false>;
    const size_t fixed = (size_t)CW * NTW * 128 * 4 + 1024;
    const size_t stage = (size_t)p.n_pad * APITCH + (size_t)TS * WSLOT;
    int stages = (int)((WIDE_SMEM_MAX - fixed) / (stage + 16));
    if (stages > 8) stages = 8;
    if (knobs().mma_stages > 0) stages = std::min(stages, std::max(2, knobs().mma_stages));
    if (stages < 2) return cudaErrorNotSupported;
    p.stages = stages;
    p.passes = (p.tiles + TS - 1) / TS;
    const size_t smem = (size_t)stages * stage + fixed + (size_t)stages * 16;
    if (const cudaError_t ea = ensure_max_dynamic_smem((const void *)kern, (int)WIDE_SMEM_MAX)) return ea;
    int64_t grid = (knobs().grid_balance & 2) ? balanced_grid(p.passes, sm_count()) : sm_count();
    if (grid > p.passes) grid = p.passes;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((CW + 1) * 32);
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

}  // namespace

size_t wide_workspace_bytes(int64_t K, int64_t N) {
    const int64_t cols = std::min<int64_t>(N, WIDE_MAX_COLS);
    const int64_t rounds = (K / QK_K + RSB - 1) / RSB;
    return (size_t)(cols > 16 ? 32 : (cols + 7) / 8 * 8) * (size_t)rounds * APITCH;
}

// F32 activations: quantizer (staged layout) + the batched decode kernel, column groups of up to 32.
// dump: parity instantiation.  cudaErrorNotSupported when a group's ring does not fit shared memory.
cudaError_t launch_gemv_q4k_wide(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x, int64_t N,
                                 int64_t x_row_stride_bytes, float *dst, int64_t dst_col_stride_bytes, void *workspace, bool pdl,
                                 cudaStream_t stream, const IntDump *dump) {
    if (M <= 0 || N <= 0) return cudaSuccess;
    const int nb = (int)(K / QK_K);
    WideParams p;
    if (!make_row_box_map(&p.wmap, w, M, (int64_t)nb * BLK_Q4K, w_row_stride_bytes, WPITCH, WROWS)) return cudaErrorNotSupported;
    for (int64_t n0 = 0; n0 < N; n0 += WIDE_MAX_COLS) {
        const int nn = (int)std::min<int64_t>(WIDE_MAX_COLS, N - n0);
        // 17..24 columns run as 32 (zero columns): measured no slower than a ragged second half, and one shape less
        const int n_pad = nn > 16 ? 32 : (nn + 7) / 8 * 8;
        const float *xs = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + n0 * x_row_stride_bytes);
        {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)(((int64_t)n_pad * nb + 7) / 8));
            cfg.blockDim = dim3(256);
            cfg.stream = stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = pdl ? 1 : 0;
            const cudaError_t eq = cudaLaunchKernelEx(&cfg, quantize_q8_K_wide_kernel, xs, (uint8_t *)workspace, nb, nn, n_pad, x_row_stride_bytes);
            count_launch();
            if (eq != cudaSuccess) return eq;
        }
        p.M = M; p.nb = nb;
        p.xs = (const uint8_t *)workspace;
        p.dst = reinterpret_cast<float *>(reinterpret_cast<char *>(dst) + n0 * dst_col_stride_bytes);
        p.dst_stride = dst_col_stride_bytes / 4;
        p.N = nn; p.n_pad = n_pad;
        p.tiles = (M + WROWS - 1) / WROWS;
        p.passes = 0; p.stages = 0;
        p.dbg = knobs().wide_dbg;
        p.dump_p = dump ? dump->p + n0 * M * nb * 8 : nullptr;
        p.dump_sumi = dump ? dump->sumi + n0 * M * nb : nullptr;
        p.dump_summ = dump ? dump->summ + n0 * M * nb : nullptr;
        cudaError_t e;
        switch (n_pad / 8) {
            // eight columns: four tiles per pass, unless that leaves SMs without a pass (M < 9472): then two tiles, eight warps
            case 1: e = p.tiles >= 4 * (int64_t)sm_count() ? launch_wide_cfg<4, 1, 1, false>(p, pdl, stream)
                                                           : launch_wide_cfg<2, 1, 1, false>(p, pdl, stream); break;
            case 2: e = launch_wide_cfg<2, 2, 1, false>(p, pdl, stream); break;
            default: e = launch_wide_cfg<2, 2, 2, false>(p, pdl, stream); break;
        }
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace q4k
