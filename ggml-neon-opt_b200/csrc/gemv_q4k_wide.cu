// gemv_q4k_wide.cu — batched decode (8 < N <= 32 columns per launch; wider batches in column groups of 32), integer dot
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
//   * a round is 4 superblocks of K.  Per round the producer issues THREE copies on mbarriers: two TMA boxes (one per tile
//     slot: 16 rows x 608 B of raw block_q4_K through a 2-D tensor map over the weight matrix — the 32 bytes beyond the
//     round's 576 are the row padding the fragment loads need, rows past M and bytes past the row read as zero) and one bulk
//     copy of the round's activation stage, which the quantizer has laid out in global memory as the exact shared-memory
//     image.  (Measured: an SM retires one bulk copy per ~80 cycles whatever its size — 48 copies of 576-1152 B per round held
//     the first version of this kernel to 2.1 TB/s of weights alone.)
//   * eight consumer warps = 4 superblocks x 2 tiles: warp w takes superblock w % 4 of tile slot w / 4.  A warp-step is
//     mma.sync.m16n8k32 (u8 x s8 -> s32) over NT = N_pad / 8 column tiles with the A fragments (weights) loaded once: the
//     "columns are batch columns" form of gemv_q4k_mma.cu, C[row][column] = the sub-block dot product p[j].
//   * a CTA works through PAIRS of tiles (a "pass"): all rounds of K for the pair, then the eight warps' partial sums meet in
//     shared memory, are added in warp order and stored.  Activations are re-streamed once per pass (from L2: the staged copy
//     of 32 columns x K = 14336 is 516 KB), weights cross HBM once.
// Staged activation block (288 B, 16-byte aligned): 256 int8 | 8 x int16 pair sums (bsums[2j] + bsums[2j+1]) | float d | pad.
// Shared-memory pitches: activation column 4 x 288 + 32 B and weight row 576 + 32 B (both 8 words mod 32): the 8-byte
// fragment loads of a half-warp (4 rows or columns x 4 pieces) hit 16 distinct bank pairs.
//
// Integers are exact (same sumi / summ as the reference: DUMP instantiation, q4k_mul_mat_dump).  Float association: per
// (row, column) every warp adds its superblocks (kb = s, s + 4, ...) in K order as  acc += (d * dy) * sumi,
// accm += (-dmin * dy) * summ;  the result is the four superblock classes' (acc + accm) added in class order.  Independent of
// grid, ring depths and of N.  Within 1e-5 of ggml_vec_dot_q4_K_q8_K (tests/test_gpu_wide.py).
#include <algorithm>

#include <cuda.h>

#include "gemv_stage.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

constexpr int WROWS = 16;                       // rows per tile (the MMA's m)
constexpr int RSB = 4;                          // superblocks per round
constexpr int TSLOTS = 2;                       // tiles per pass
constexpr int CW = RSB * TSLOTS;                // consumer warps
constexpr int WIDE_THREADS = (CW + 1) * 32;     // + the producer warp
constexpr int XBLK = 288;                       // staged activation block
constexpr int XD_OFF = 272;                     // float d inside it (pair sums at 256)
constexpr int APITCH = RSB * XBLK + 32;         // activation column pitch in a stage (1184 B = 296 words = 8 mod 32)
constexpr int WPITCH = RSB * BLK_Q4K + 32;      // weight row pitch in a slot (608 B = 152 words = 24 mod 32)
constexpr int WSLOT = WROWS * WPITCH;           // 9728 B
constexpr size_t WIDE_SMEM_MAX = 232448;

struct WideParams {
    CUtensorMap wmap;           // the weights as a 2-D tensor of 8-byte elements [M][row bytes / 8], box = 16 rows x 608 B
    const uint8_t *w;
    int64_t M, w_stride;
    int nb;
    const uint8_t *xs;          // staged activations: [rounds][n_pad] column pieces of APITCH bytes (4 blocks of XBLK + pad) = the stage images
    float *dst;
    int64_t dst_stride;         // floats between result columns
    int N, n_pad;
    int64_t tiles, passes;      // 16-row tiles, tile pairs
    int ast, wst;               // ring depths: activation stages, weight rounds (TSLOTS slots each)
    int dbg;                    // development (knob wide_dbg): 1 = no activation copies, 2 = no arithmetic, 4 = no weight copies — wrong results, timing only
    int32_t *dump_p, *dump_sumi, *dump_summ;
};

__device__ __forceinline__ void imma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void tma_box_g2s(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ uint2 lds64w(const uint8_t *p) { return *reinterpret_cast<const uint2 *>(p); }
__device__ __forceinline__ void bar_sync_named(int id, int threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory"); }

// summ of one (row, column): sum_j m[j] * ps[j]; mz / mw = the row's eight mins as bytes, ps = the column's pair sums
__device__ __forceinline__ int mins_dot(const uint4 ps, uint32_t mz, uint32_t mw) {
    return __dp2a_hi((int)ps.w, (int)mw, __dp2a_lo((int)ps.z, (int)mw, __dp2a_hi((int)ps.y, (int)mz, __dp2a_lo((int)ps.x, (int)mz, 0))));
}

template <int NT, bool DUMP>
__global__ void __launch_bounds__(WIDE_THREADS, 1)
gemv_q4k_wide_kernel(const __grid_constant__ WideParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nb = p.nb, AST = p.ast, WST = p.wst;
    const int rounds = (nb + RSB - 1) / RSB;
    const size_t astage = (size_t)p.n_pad * APITCH;

    size_t off = 0;
    uint8_t *act_s = smem;                                      off += (size_t)AST * astage;
    uint8_t *w_s = smem + off;                                  off += (size_t)WST * TSLOTS * WSLOT;
    float *red_s = reinterpret_cast<float *>(smem + off);       off += (size_t)CW * NT * 128 * 4;       // [warp][nt][i][lane]
    uint64_t *afull = reinterpret_cast<uint64_t *>(smem + off); off += (size_t)AST * 8;
    uint64_t *aempty = reinterpret_cast<uint64_t *>(smem + off); off += (size_t)AST * 8;
    uint64_t *wfull = reinterpret_cast<uint64_t *>(smem + off); off += (size_t)WST * TSLOTS * 8;
    uint64_t *wempty = reinterpret_cast<uint64_t *>(smem + off); off += (size_t)WST * TSLOTS * 8;

    if (threadIdx.x == 0) {
        for (int s = 0; s < AST; ++s) { mbar_init(smem_u32(&afull[s]), 1); mbar_init(smem_u32(&aempty[s]), CW); }
        for (int s = 0; s < WST * TSLOTS; ++s) { mbar_init(smem_u32(&wfull[s]), 1); mbar_init(smem_u32(&wempty[s]), RSB); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    griddep_launch_dependents();

    // passes of this CTA: tile pairs blockIdx.x, blockIdx.x + gridDim.x, ... (the grid sweeps the matrix front to back)
    const int64_t my_passes = p.passes > blockIdx.x ? (p.passes - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total = my_passes * rounds;                  // rounds this CTA walks through, numbered q = 0 .. total - 1

    if (warp == CW) {
        // ---- producer warp ----
        // running state of the two copy sequences (no 64-bit divisions by run-time values on the issue path)
        struct Seq { int64_t pass; int rd, slot; uint32_t ph; };
        Seq sw{(int64_t)blockIdx.x, 0, 0, 0u}, sa{(int64_t)blockIdx.x, 0, 0, 0u};
        auto advance = [&](Seq &s, int depth) {
            if (++s.rd == rounds) { s.rd = 0; s.pass += gridDim.x; }
            if (++s.slot == depth) { s.slot = 0; s.ph ^= 1u; }
        };
        auto issue_w = [&]() {                                 // lanes 0 and 1: one TMA box each (tile slots 0 and 1)
            if (lane < TSLOTS) {
                const int ts = lane;
                mbar_wait(smem_u32(&wempty[sw.slot * TSLOTS + ts]), sw.ph ^ 1u);
                const uint32_t bar = smem_u32(&wfull[sw.slot * TSLOTS + ts]);
                if (p.dbg & 4) mbar_arrive(bar);
                else {
                    mbar_expect_tx(bar, WSLOT);
                    fence_proxy_async();
                    tma_box_g2s(smem_u32(w_s + (size_t)(sw.slot * TSLOTS + ts) * WSLOT), &p.wmap, sw.rd * (RSB * BLK_Q4K / 8),
                                (int)((sw.pass * TSLOTS + ts) * WROWS), bar);
                }
            }
            advance(sw, WST);
        };
        auto issue_a = [&]() {
            if (lane == 0) {
                mbar_wait(smem_u32(&aempty[sa.slot]), sa.ph ^ 1u);
                const uint32_t bar = smem_u32(&afull[sa.slot]);
                if (p.dbg & 1) mbar_arrive(bar);
                else {
                    mbar_expect_tx(bar, (uint32_t)astage);
                    fence_proxy_async();
                    bulk_g2s(smem_u32(act_s + (size_t)sa.slot * astage), p.xs + (size_t)sa.rd * astage, (uint32_t)astage, bar);
                }
            }
            advance(sa, AST);
        };
        // weights never depend on the previous kernel in the stream (the quantizer): the first WST rounds go out before the wait
        // (both rings refill on the same event — round q consumed — so the loop waits in consumption order: a wait for a
        // later round in front of a copy for an earlier one would cut that ring's lookahead to one round)
        for (int64_t q = 0; q < WST && q < total; ++q) issue_w();
        griddep_wait();
        for (int64_t q = 0; q < AST && q < total; ++q) issue_a();
        for (int64_t q = 0; q < total; ++q) {
            if (q + AST < total) issue_a();
            if (q + WST < total) issue_w();
        }
        return;
    }

    // ---- consumer warps ----
    const int sb = warp & (RSB - 1), ts = warp >> 2;
    const int r = lane >> 2, c = lane & 3;
    int ast = 0, wsr = 0;                                       // ring positions and their phases, advanced once per round
    uint32_t aph = 0u, wph = 0u;
    for (int64_t ip = 0; ip < my_passes; ++ip) {
        const int64_t pass = blockIdx.x + ip * gridDim.x;
        const int64_t tile = pass * TSLOTS + ts;
        const bool tile_ok = tile < p.tiles;
        float acc[NT][4], accm[NT][4];
#pragma unroll
        for (int t = 0; t < NT; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) { acc[t][i] = 0.0f; accm[t][i] = 0.0f; }

        for (int rd = 0; rd < rounds; ++rd) {
            const int wsl = wsr * TSLOTS + ts;
            mbar_wait(smem_u32(&afull[ast]), aph);
            mbar_wait(smem_u32(&wfull[wsl]), wph);
            const int kb = rd * RSB + sb;
            if (kb < nb && tile_ok && !(p.dbg & 2)) {
                const uint8_t *p0 = w_s + (size_t)wsl * WSLOT + r * WPITCH + sb * BLK_Q4K, *p1 = p0 + 8 * WPITCH;
                const uint8_t *ya = act_s + (size_t)ast * astage + sb * XBLK;
                const uint4 h0 = lds128(p0), h1 = lds128(p1);
                const uint4 u0 = unpack_scales_mins(h0), u1 = unpack_scales_mins(h1);
                int s[NT][4];
#pragma unroll
                for (int t = 0; t < NT; ++t) { s[t][0] = 0; s[t][1] = 0; s[t][2] = 0; s[t][3] = 0; }
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const uint2 wq = lds64w(p0 + 16 + 32 * g + 8 * c), vq = lds64w(p1 + 16 + 32 * g + 8 * c);
                    const uint32_t a0 = wq.x & 0x0f0f0f0fu, a1 = vq.x & 0x0f0f0f0fu, a2 = wq.y & 0x0f0f0f0fu, a3 = vq.y & 0x0f0f0f0fu;
                    const uint32_t e0 = wq.x & 0xf0f0f0f0u, e1 = vq.x & 0xf0f0f0f0u, e2 = wq.y & 0xf0f0f0f0u, e3 = vq.y & 0xf0f0f0f0u;
                    const uint32_t q0 = (g < 2 ? u0.x : u0.y) >> (16 * (g & 1)), q1 = (g < 2 ? u1.x : u1.y) >> (16 * (g & 1));
                    const int lo0 = (int)((q0 & 0xffu) << 4), hi0 = (int)((q0 >> 8) & 0xffu);
                    const int lo1 = (int)((q1 & 0xffu) << 4), hi1 = (int)((q1 >> 8) & 0xffu);
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        const uint8_t *ycol = ya + (size_t)(8 * t + r) * APITCH + 64 * g + 8 * c;     // this lane's B column = 8t + lane / 4
                        const uint2 bl = lds64w(ycol), bh = lds64w(ycol + 32);
                        int cl[4] = {0, 0, 0, 0}, ch[4] = {0, 0, 0, 0};
                        imma(cl, a0, a1, a2, a3, bl.x, bl.y);
                        imma(ch, e0, e1, e2, e3, bh.x, bh.y);
                        s[t][0] += lo0 * cl[0] + hi0 * ch[0];
                        s[t][1] += lo0 * cl[1] + hi0 * ch[1];
                        s[t][2] += lo1 * cl[2] + hi1 * ch[2];
                        s[t][3] += lo1 * cl[3] + hi1 * ch[3];
                        if constexpr (DUMP) {      // C[row][column] of this MMA pair IS p[2g] (and 16 x p[2g + 1]) of that (row, column)
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const int64_t row = tile * WROWS + r + 8 * (i >> 1);
                                const int col = 8 * t + 2 * c + (i & 1);
                                if (row < p.M && col < p.N) {
                                    int32_t *o = p.dump_p + (((int64_t)col * p.M + row) * nb + kb) * 8 + 2 * g;
                                    o[0] = cl[i]; o[1] = ch[i] >> 4;
                                }
                            }
                        }
                    }
                }
                const float2 dd0 = __half22float2(*reinterpret_cast<const __half2 *>(&h0.x));     // {d, dmin} of row r
                const float2 dd1 = __half22float2(*reinterpret_cast<const __half2 *>(&h1.x));     // row r + 8
#pragma unroll
                for (int t = 0; t < NT; ++t) {
                    const uint8_t *ylo = ya + (size_t)(8 * t + 2 * c) * APITCH, *yhi = ylo + APITCH;   // columns 8t + 2c, 8t + 2c + 1
                    const uint4 pl = lds128(ylo + 256), ph = lds128(yhi + 256);
                    const float yd_lo = *reinterpret_cast<const float *>(ylo + XD_OFF), yd_hi = *reinterpret_cast<const float *>(yhi + XD_OFF);
                    const int ml0 = mins_dot(pl, u0.z, u0.w), mh0 = mins_dot(ph, u0.z, u0.w);
                    const int ml1 = mins_dot(pl, u1.z, u1.w), mh1 = mins_dot(ph, u1.z, u1.w);
                    if constexpr (DUMP) {
                        const int si[4] = {s[t][0] >> 4, s[t][1] >> 4, s[t][2] >> 4, s[t][3] >> 4}, sm[4] = {ml0, mh0, ml1, mh1};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int64_t row = tile * WROWS + r + 8 * (i >> 1);
                            const int col = 8 * t + 2 * c + (i & 1);
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
            __syncwarp();
            if (lane == 0) { mbar_arrive(smem_u32(&aempty[ast])); mbar_arrive(smem_u32(&wempty[wsl])); }
            if (++ast == AST) { ast = 0; aph ^= 1u; }
            if (++wsr == WST) { wsr = 0; wph ^= 1u; }
        }

        // ---- end of the pass: the four superblock classes of each tile meet, in class order ----
        float *mine = red_s + (size_t)warp * NT * 128;
#pragma unroll
        for (int t = 0; t < NT; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) mine[(t * 4 + i) * 32 + lane] = __fadd_rn(acc[t][i], accm[t][i]);
        bar_sync_named(1, CW * 32);
        if (sb == 0 && tile_ok) {                                  // warps 0 and 4: one per tile slot
#pragma unroll
            for (int t = 0; t < NT; ++t) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float *pp = red_s + (size_t)(ts * RSB) * NT * 128 + (t * 4 + i) * 32 + lane;
                    float a = pp[0];
#pragma unroll
                    for (int k = 1; k < RSB; ++k) a = __fadd_rn(a, pp[(size_t)k * NT * 128]);
                    const int64_t row = tile * WROWS + r + 8 * (i >> 1);
                    const int col = 8 * t + 2 * c + (i & 1);
                    if (row < p.M && col < p.N) p.dst[(int64_t)col * p.dst_stride + row] = a;
                }
            }
        }
        bar_sync_named(1, CW * 32);                                // red_s is free again
    }
}

// Q8_K quantizer into the staged layout — the shared-memory image of every round's activation stage, one after the other:
// [round][column] pieces of APITCH bytes (the round's 4 blocks of the column + 32 B of bank padding).  One warp per
// (column, superblock); columns >= N (padding up to a multiple of 8) are written as zero blocks.  Same arithmetic as quantize_q8k.cu (quantize_block_q8_K: explicit single-rounding intrinsics).
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
        const float4 a = reinterpret_cast<const float4 *>(src)[lane], e = reinterpret_cast<const float4 *>(src)[32 + lane];
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

template <int NT>
cudaError_t launch_wide_nt(WideParams p, bool pdl, cudaStream_t stream) {
    auto kern = p.dump_sumi ? gemv_q4k_wide_kernel<NT, true> : gemv_q4k_wide_kernel<NT, false>;
    const size_t fixed = (size_t)CW * NT * 128 * 4 + 1024;
    const size_t astage = (size_t)p.n_pad * APITCH, wround = (size_t)TSLOTS * WSLOT;
    // ring depths: at least 2 activation stages and 3 weight rounds; what is left goes to whichever is cheaper per round of lookahead
    int ast = 2, wst = 3;
    while (fixed + (size_t)(ast + 1) * astage + (size_t)(wst + 1) * wround + (size_t)(ast + wst + 2) * 48 <= WIDE_SMEM_MAX && ast < 4) { ++ast; ++wst; }
    while (fixed + (size_t)ast * astage + (size_t)(wst + 1) * wround + (size_t)(ast + wst + 1) * 48 <= WIDE_SMEM_MAX && wst < 8) ++wst;
    if (knobs().mma_stages > 0) { ast = std::min(ast, std::max(2, knobs().mma_stages)); wst = std::min(wst, std::max(2, knobs().mma_stages)); }
    p.ast = ast; p.wst = wst;
    const size_t smem = fixed + (size_t)ast * astage + (size_t)wst * wround + (size_t)(ast + wst * TSLOTS) * 16;
    if (smem > WIDE_SMEM_MAX) return cudaErrorNotSupported;
    if (const cudaError_t ea = ensure_max_dynamic_smem((const void *)kern, (int)WIDE_SMEM_MAX)) return ea;
    int64_t grid = sm_count();
    if (grid > p.passes) grid = p.passes;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(WIDE_THREADS);
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
    return (size_t)((cols + 7) / 8 * 8) * (size_t)rounds * APITCH;
}

namespace {
// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda)
cudaError_t make_wide_weight_map(CUtensorMap *map, const void *w, int64_t M, int64_t nb, int64_t w_row_stride_bytes) {
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        const cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess) return e;
        if (qres != cudaDriverEntryPointSuccess || !fn) return cudaErrorNotSupported;
        encode = (encode_fn)fn;
    }
    const cuuint64_t dims[2] = {(cuuint64_t)(nb * BLK_Q4K / 8), (cuuint64_t)M};
    const cuuint64_t strides[1] = {(cuuint64_t)w_row_stride_bytes};
    const cuuint32_t box[2] = {(cuuint32_t)(WPITCH / 8), (cuuint32_t)WROWS}, estr[2] = {1, 1};
    const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_INT64, 2, const_cast<void *>(w), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}
}  // namespace

// F32 activations, 8 < N: quantizer (staged layout) + the batched decode kernel, column groups of up to 32.
// dump: parity instantiation (p may be null).  cudaErrorNotSupported when a group's rings do not fit shared memory.
cudaError_t launch_gemv_q4k_wide(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x, int64_t N,
                                 int64_t x_row_stride_bytes, float *dst, int64_t dst_col_stride_bytes, void *workspace, bool pdl,
                                 cudaStream_t stream, const IntDump *dump) {
    if (M <= 0 || N <= 0) return cudaSuccess;
    const int nb = (int)(K / QK_K);
    for (int64_t n0 = 0; n0 < N; n0 += WIDE_MAX_COLS) {
        const int nn = (int)std::min<int64_t>(WIDE_MAX_COLS, N - n0);
        const int n_pad = (nn + 7) / 8 * 8;
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
        WideParams p;
        if (const cudaError_t em = make_wide_weight_map(&p.wmap, w, M, nb, w_row_stride_bytes)) return em;
        p.w = (const uint8_t *)w; p.M = M; p.w_stride = w_row_stride_bytes; p.nb = nb;
        p.xs = (const uint8_t *)workspace;
        p.dst = reinterpret_cast<float *>(reinterpret_cast<char *>(dst) + n0 * dst_col_stride_bytes);
        p.dst_stride = dst_col_stride_bytes / 4;
        p.N = nn; p.n_pad = n_pad;
        p.tiles = (M + WROWS - 1) / WROWS;
        p.passes = (p.tiles + TSLOTS - 1) / TSLOTS;
        p.ast = p.wst = 0;
        p.dbg = knobs().wide_dbg;
        p.dump_p = dump ? dump->p + n0 * M * nb * 8 : nullptr;
        p.dump_sumi = dump ? dump->sumi + n0 * M * nb : nullptr;
        p.dump_summ = dump ? dump->summ + n0 * M * nb : nullptr;
        cudaError_t e;
        switch (n_pad / 8) {
            case 1: e = launch_wide_nt<1>(p, pdl, stream); break;
            case 2: e = launch_wide_nt<2>(p, pdl, stream); break;
            case 3: e = launch_wide_nt<3>(p, pdl, stream); break;
            default: e = launch_wide_nt<4>(p, pdl, stream); break;
        }
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace q4k
