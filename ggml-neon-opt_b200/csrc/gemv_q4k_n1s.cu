// gemv_q4k_n1s.cu — single-column decode (N == 1) for matrices that stream tens of megabytes or have a long K: the data path of
// the batched decode kernel (gemv_q4k_wide.cu) under the arithmetic of the tensor-pipe decode kernel (gemv_q4k_mma.cu, N == 1).
// Replaces the vec_dot loop of ggml_compute_forward_mul_mat_one_chunk (L/ggml/src/ggml-cpu/ggml-cpu.c:1115-1203) +
// ggml_vec_dot_q4_K_q8_K (L/ggml/src/ggml-cpu/arch/x86/quants.c:1741-1823) at N == 1.
//
// Why a third N == 1 launch shape.  gemv_q4k_mma.cu and gemv_q4k_tma.cu hold the whole Q8_K activation vector in every CTA's
// shared memory: every CTA quantizes it (K <= 16384) or copies it from the workspace (longer K: 112 blocks = 32 KB per CTA at
// K = 28672) before its first tile, and the memory pipe idles meanwhile — 8192 x 28672 ran at 74% of the measured HBM peak where
// 28672 x 8192 reached 83%, and 4096 x 14336 at 43%.  Here the activations are quantized ONCE (quantize_q8_K_n1s_kernel) and
// STREAM through shared memory next to the weights, 2304 bytes per round: a CTA's first weight tile is consumed as soon as the
// quantizer in front of it has finished, and nothing is staged per CTA.
//
// Data path: a round = 8 superblocks of K for two 16-row tiles.  One lane issues three copies per round on one mbarrier: two TMA
// boxes (16 rows x 1168 B of raw block_q4_K through a 2-D tensor map; rows past M / bytes past the row read as zero) and one
// bulk copy of the round's eight staged activation blocks.  Sixteen consumer warps = 2 tiles x 8 superblock classes.
//
// Arithmetic and bits: the superblock step IS gemv_q4k_mma.cu's (gemv_mma_step.cuh: sb_step_n1), warp class w takes superblock
// 8 * round + w, segments of four rounds, chunk lanes summed by xor 1 / xor 2, the eight classes added in order, segments in K
// order — the float association of gemv_q4k_mma.cu and of the large-matrix launch of gemv_q4k_tma.cu (ORD 2).  A row's bits do
// not depend on which of the three kernels computed it (tests/test_gpu_n1s.py compares them bit for bit).
#include <algorithm>

#include <cuda.h>

#include "gemv_mma_step.cuh"
#include "gemv_stage.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

constexpr int NS_TS = 2;                              // tiles per pass
constexpr int NS_CW = NS_TS * KW;                     // consumer warps
constexpr int NS_THREADS = (NS_CW + 1) * 32;          // + the producer warp
constexpr int NS_XBLK = 288;                          // staged activation block: 256 int8 | 8 x int16 pair sums | float d | pad
constexpr int NS_XD = 272;
constexpr int NS_ABYTES = KW * NS_XBLK;               // activation bytes of a round
constexpr int NS_SLOT = MmaCfg<1>::SLOT;              // 16 rows x 1168 B
constexpr int NS_STAGE = NS_TS * NS_SLOT + NS_ABYTES; // 39680 B
constexpr int NS_SEG = 4;                             // rounds per segment (gemv_q4k_mma.cu SEG_ROUNDS)
constexpr size_t NS_SMEM_MAX = 232448;

struct N1sParams {
    CUtensorMap wmap;
    int64_t M;
    int nb;
    const uint8_t *xs;          // nb staged activation blocks
    float *dst;
    int64_t tiles, passes;
    int stages;
    // LL exchange of the row-sharded path (q4k_kernels.h LLExchange; the field names gemv_stage.cuh's ll_* helpers use):
    // ll_world > 0 = row `dst_row0 + m` leaves as a {value, epoch} pair for every rank and the tail of the launch gathers
    uint2 *ll_peers[8];
    const uint2 *ll_local;
    int64_t ll_half, ll_M, ll_dst_stride;
    uint32_t *ll_cnt;
    float *ll_dst;
    int ll_world, ll_rank;
    int64_t dst_row0;
};

template <bool LL>
__global__ void __launch_bounds__(NS_THREADS, 1)
gemv_q4k_n1s_kernel(const __grid_constant__ N1sParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nb = p.nb, S = p.stages;
    const int rounds = (nb + KW - 1) / KW, nseg = (rounds + NS_SEG - 1) / NS_SEG;

    uint8_t *ring = smem;
    uint8_t *zero_s = smem + (size_t)S * NS_STAGE;                                    // one zero block (B columns that carry nothing)
    float *part_s = reinterpret_cast<float *>(zero_s + NS_XBLK);                      // [2 buffers][tile slot][class][16 rows]
    uint64_t *full = reinterpret_cast<uint64_t *>(part_s + 2 * NS_TS * KW * MROWS);
    uint64_t *empty = full + S;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full[s]), 1); mbar_init(smem_u32(&empty[s]), NS_CW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < NS_XBLK / 4) reinterpret_cast<uint32_t *>(zero_s)[threadIdx.x] = 0u;
    __syncthreads();
    griddep_launch_dependents();

    const int64_t my_passes = p.passes > blockIdx.x ? (p.passes - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total = my_passes * rounds;

    if (warp == NS_CW) {
        // ---- producer: one lane, three copies per round ----
        if (lane != 0) return;
        int64_t pass = blockIdx.x;
        int rd = 0, slot = 0;
        uint32_t ph = 0u;
        bool waited = false;                                   // the activations depend on the quantizer in front of this kernel
        for (int64_t q = 0; q < total; ++q) {
            mbar_wait(smem_u32(&empty[slot]), ph ^ 1u);
            const uint32_t bar = smem_u32(&full[slot]);
            uint8_t *st = ring + (size_t)slot * NS_STAGE;
            mbar_expect_tx(bar, NS_STAGE);
            fence_proxy_async();
#pragma unroll
            for (int ts = 0; ts < NS_TS; ++ts)
                tma_box_g2s(smem_u32(st + (size_t)ts * NS_SLOT), &p.wmap, rd * (ROW_BYTES / 8), (int)((pass * NS_TS + ts) * MROWS), bar);
            if (!waited) { griddep_wait(); waited = true; }
            // (the last round may reach past the staged blocks of a ragged K: the workspace is sized in whole rounds)
            bulk_g2s(smem_u32(st + NS_TS * NS_SLOT), p.xs + (size_t)rd * NS_ABYTES, NS_ABYTES, bar);
            if (++rd == rounds) { rd = 0; pass += gridDim.x; }
            if (++slot == S) { slot = 0; ph ^= 1u; }
        }
        return;
    }

    // ---- consumer warps: (tile slot, superblock class) ----
    const int cls = warp & (KW - 1), ts = warp >> 3;
    const int r = lane >> 2, c = lane & 3;
    int slot = 0, buf = 0;
    uint32_t ph = 0u;
    const DumpAt da{};
    uint32_t ll_ep = 0u;
    if (LL && cls == 0) {       // the storing warps: this call's epoch (the previous call of the object has completed behind the wait)
        griddep_wait();
        ll_ep = ll_epoch(p);
        if (blockIdx.x == 0 && ts == 0) ll_presence(p, ll_ep, lane);
    }
    for (int64_t ip = 0; ip < my_passes; ++ip) {
        const int64_t tile = (blockIdx.x + ip * gridDim.x) * NS_TS + ts;
        const bool tile_ok = tile < p.tiles;
        float tot = 0.0f;                                       // class-0 warp, lane = row of the tile
        for (int seg = 0; seg < nseg; ++seg) {
            float acc[2] = {0.0f, 0.0f}, accm[2] = {0.0f, 0.0f};
            const int rde = min((seg + 1) * NS_SEG, rounds);
            for (int rd = seg * NS_SEG; rd < rde; ++rd) {
                mbar_wait(smem_u32(&full[slot]), ph);
                const int kb = rd * KW + cls;
                if (kb < nb && tile_ok) {
                    const uint8_t *st = ring + (size_t)slot * NS_STAGE;
                    const uint8_t *yblk = st + NS_TS * NS_SLOT + cls * NS_XBLK;
                    const int j = r;                                     // this lane's B column = sub-block j
                    const uint8_t *y0 = (j == 2 * c) ? yblk + 32 * j : zero_s;
                    const uint8_t *y1 = (j == 2 * c + 1) ? yblk + 32 * j : zero_s;
                    sb_step_n1<false>(st + (size_t)ts * NS_SLOT + cls * BLK_Q4K, y0, y1, yblk, *reinterpret_cast<const float *>(yblk + NS_XD), lane,
                                      acc, accm, da);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&empty[slot]));
                if (++slot == S) { slot = 0; ph ^= 1u; }
            }
            // segment epilogue: chunk lanes, then the eight classes in order (gemv_q4k_mma.cu's consumer / reducer pair)
            float v0 = __fadd_rn(acc[0], accm[0]), v1 = __fadd_rn(acc[1], accm[1]);
            v0 = __fadd_rn(v0, __shfl_xor_sync(0xffffffffu, v0, 1));
            v1 = __fadd_rn(v1, __shfl_xor_sync(0xffffffffu, v1, 1));
            v0 = __fadd_rn(v0, __shfl_xor_sync(0xffffffffu, v0, 2));
            v1 = __fadd_rn(v1, __shfl_xor_sync(0xffffffffu, v1, 2));
            float *part = part_s + ((size_t)(buf * NS_TS + ts) * KW) * MROWS;
            if (c == 0) { part[cls * MROWS + r] = v0; part[cls * MROWS + r + 8] = v1; }
            named_bar_sync(1 + ts, KW * 32);                     // the eight warps of this tile slot
            if (cls == 0 && lane < MROWS) {
                float a = part[lane];
#pragma unroll
                for (int w = 1; w < KW; ++w) a = __fadd_rn(a, part[w * MROWS + lane]);
                tot = seg == 0 ? a : __fadd_rn(tot, a);
            }
            buf ^= 1;           // the other buffer next: this one is read before its readers arrive at the next barrier
        }
        if (cls == 0 && lane < MROWS && tile_ok) {
            const int64_t row = tile * MROWS + lane;
            if (row < p.M) {
                if (LL) ll_store(p, ll_ep, 0, p.dst_row0 + row, tot);
                else p.dst[row] = tot;
            }
        }
    }
    // the gather: one warp per CTA takes its slice of the full vector out of this rank's LL buffer (every CTA is resident)
    if (LL && warp == 0) ll_tail(p, 1, ll_ep, blockIdx.x, gridDim.x, lane, 32);
}

// quantize_row_q8_K for one activation row into staged blocks (the arithmetic of quantize_q8k.cu; one warp per block)
__global__ void __launch_bounds__(256)
quantize_q8_K_n1s_kernel(const float *__restrict__ x, uint8_t *__restrict__ xs, int nb) {
    griddep_launch_dependents();
    griddep_wait();
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (b >= nb) return;
    const float *src = x + (int64_t)b * QK_K;
    const float4 a = ld_act16(src + 4 * lane), e = ld_act16(src + 128 + 4 * lane);
    const float v[8] = {a.x, a.y, a.z, a.w, e.x, e.y, e.z, e.w};
    uint32_t w_lo, w_hi;
    int s_lo, s_hi;
    const float d = quantize_block_q8_K<8>(v, lane, w_lo, w_hi, s_lo, s_hi);
    uint8_t *dst = xs + (size_t)b * NS_XBLK;
    reinterpret_cast<uint32_t *>(dst)[lane] = w_lo;
    reinterpret_cast<uint32_t *>(dst)[32 + lane] = w_hi;
    if ((lane & 7) == 0) {
        reinterpret_cast<int16_t *>(dst + 256)[lane >> 3] = (int16_t)s_lo;
        reinterpret_cast<int16_t *>(dst + 256)[4 + (lane >> 3)] = (int16_t)s_hi;
    }
    if (lane == 0) *reinterpret_cast<float *>(dst + NS_XD) = d;
    if (lane >= 1 && lane <= 3) reinterpret_cast<uint32_t *>(dst + NS_XD)[lane] = 0u;
}

}  // namespace

size_t n1s_workspace_bytes(int64_t K) { return (size_t)((K / QK_K + KW - 1) / KW) * NS_ABYTES; }

// F32 activation row -> quantizer + the streamed single-column kernel.  cudaErrorNotSupported when no tensor map can be made.
cudaError_t launch_gemv_q4k_n1s(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x, float *dst,
                                void *workspace, bool pdl, cudaStream_t stream, const LLExchange *ll, int64_t dst_row0) {
    if (M <= 0) return cudaSuccess;
    const int nb = (int)(K / QK_K);
    N1sParams p;
    if (!make_row_box_map(&p.wmap, w, M, (int64_t)nb * BLK_Q4K, w_row_stride_bytes, MmaCfg<1>::PITCH, MROWS)) return cudaErrorNotSupported;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)((nb + 7) / 8));
        cfg.blockDim = dim3(256);
        cfg.stream = stream;
        cfg.attrs = attr;
        cfg.numAttrs = pdl ? 1 : 0;
        const cudaError_t eq = cudaLaunchKernelEx(&cfg, quantize_q8_K_n1s_kernel, x, (uint8_t *)workspace, nb);
        count_launch();
        if (eq != cudaSuccess) return eq;
    }
    p.M = M; p.nb = nb;
    p.xs = (const uint8_t *)workspace;
    p.dst = dst;
    fill_ll(p, ll);
    p.dst_row0 = dst_row0;
    auto kern = ll ? gemv_q4k_n1s_kernel<true> : gemv_q4k_n1s_kernel<false>;
    p.tiles = (M + MROWS - 1) / MROWS;
    p.passes = (p.tiles + NS_TS - 1) / NS_TS;
    const size_t fixed = NS_XBLK + (size_t)2 * NS_TS * KW * MROWS * 4 + 256;
    int stages = (int)((NS_SMEM_MAX - fixed) / (NS_STAGE + 16));
    if (knobs().mma_stages > 0) stages = std::min(stages, std::max(2, knobs().mma_stages));
    p.stages = stages;
    const size_t smem = (size_t)stages * NS_STAGE + fixed + (size_t)stages * 16;
    if (const cudaError_t ea = ensure_max_dynamic_smem((const void *)kern, (int)NS_SMEM_MAX)) return ea;
    // every CTA the same number of passes: 8192 x 28672 is 256 passes — on 148 CTAs the second round runs with 108 of them, below
    // what the memory system can feed (26.6 us); on 128 CTAs both rounds run full (24.0 us = 85% of the measured HBM peak)
    int64_t grid = (knobs().grid_balance & 1) ? balanced_grid(p.passes, sm_count()) : sm_count();
    if (knobs().n1s_grid > 0) grid = std::min<int64_t>(grid, knobs().n1s_grid);
    if (grid > p.passes) grid = p.passes;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(NS_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
    count_launch();
    return e;
}

}  // namespace q4k
