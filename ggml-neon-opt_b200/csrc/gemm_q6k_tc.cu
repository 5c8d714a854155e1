// gemm_q6k_tc.cu — Q6_K prefill GEMM (batch > 32) on the 5th-generation tensor cores: tcgen05.mma kind::f16, TMEM accumulators,
// weights fed to the tensor core from TENSOR MEMORY, the skeleton of gemm_q4k_tc.cu.
//   dst[n][m] = sum_blk  d_x[n,blk] * ( d_w[m,blk] * sumi[m,n,blk] )
// which is ggml_compute_forward_mul_mat for Q6_K x F32 at ne11 = N (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401) with the integer
// stage of ggml_vec_dot_q6_K_q8_K (L/ggml/src/ggml-cpu/quants.c:795-848; AVX2: arch/x86/quants.c:2129-2218) kept EXACT:
//   sumi = sum_{sub < 16} sc[sub] * sum_{l < 16} (q6[16 sub + l] - 32) * q8[16 sub + l]
//
// Why not the int8 tensor core as for Q4_K.  a = q6 - 32 is in [-32, 31] and sc is a full signed byte, so a * sc needs 14 bits;
// any split of the SCALE whose pieces keep a * piece inside a signed byte has four pieces (|piece| <= 3), and a split of the
// PRODUCT costs shifts, masks and byte packing per ELEMENT in the producer threads, which are what bounds gemm_q4k_tc.cu.
// fp16 does it in two pieces with one packed instruction per pair of weights and piece:
//   sc = 16 sh + sl   (sl = sc & 15 in [0, 15], sh = sc >> 4 in [-8, 7])
//   A_hi = a * sh  (|.| <= 256),  A_lo = a * sl  (|.| <= 480):  small integers, exact in fp16; the activations q8 (|.| <= 127)
//   are exact in fp16 as well, every product is an integer below 2^16, and a whole superblock's sum is below
//   256 * 480 * 127 < 2^24: the fp32 accumulator of kind::f16 holds it without rounding (checked against the oracle's
//   integers by the DUMP instantiation, tests/test_gpu_q6k_gemm.py).  sumi = 16 * S_hi + S_lo; the epilogue forms
//   (float)sumi as fma(S_hi, 16, S_lo) — one rounding of the exact integer, which is what the reference's int -> float does.
//   Building the pieces: the 6-bit q lands in the low byte of an fp16 with exponent byte 0x64 (= 1024 + q, one PRMT per
//   pair), and  fma(1024 + q, s, -1056 s) = (q - 32) s  exactly (the constants 1056 s are multiples of 32 below 2^14).
// The price: kind::f16 runs at half the int8 rate and the B tile is two bytes per activation; two f16 MMAs per weight make the
// ceiling a quarter of the dense fp16 rate (~1.1 POPS of useful int ops on B200).
//
// CTA = 128 weight rows x 128 activation columns, 27 warps, warp-specialised; a pipeline step is 64 elements of K (4 per
// superblock):
//   warps 0-15   epilogue (thread <-> TMEM lane = weight row, column group g = warp / 4 of both 64-column halves) AND the
//                weight feed: Q6_K rows are 210-byte blocks at 2-byte alignment, which neither TMA nor cp.async can lay down
//                aligned, so these warps — idle most of a superblock — stage the 53 aligned words that cover each row's block
//                (4-byte cp.async) and, a superblock later, unpack the stage into the image the producers read: 256 six-bit
//                values as bytes in element order + 16 scales, row pitch 272 B (conflict-free 16-byte reads), two superblocks
//                ahead of the tensor core.
//   warps 16-23  producers: thread <-> weight row (quadrant = warp % 4) and K half (warp / 4) of the step: two 16-byte reads,
//                one PRMT and two HFMA2 per pair of weights, tcgen05.st into one of four A stages.
//   warps 24-25  allocate TMEM (24); one lane of each issues the MMAs (M = 128, N = 64, K = 16) of ITS column half and commits
//                to mbarriers.
//   warp 26      one lane: bulk copies of the activation tiles (pre-tiled by the quantizer in exactly the shared-memory order).
// TMEM columns: [0,128) S_hi, [128,256) S_lo, [256,512) four A stages x (hi, lo) x 32.
// K order inside a step (both operands agree, any order is a valid dot product): step (n, hp) of a superblock holds the 64
// elements 128 n + 32 e + 16 hp + j  (e < 4, j < 16) at slot 16 e + j — the four runs of 16 that one 16-byte piece of ql / qh
// covers, so a producer thread reads three aligned 16-byte words per step and its two scales are constant over 16 slots.
#include <algorithm>

#include <cuda_fp16.h>

#include "gemm_tc_ptx.cuh"
#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

constexpr int BLK_Q6K = 210;
constexpr int BN = 128, KS = 64, SB_STEPS = QK_K / KS;      // column tile; K elements per step; steps per superblock
constexpr int B_SLOTS = 6, A_STAGES = 4, Q_SLOTS = 2;
constexpr int TILE_BYTES = BN * KS * 2;                      // 16 KB: the fp16 activation tile of a step
constexpr int RAW_PITCH = 240, RAW_WORDS = 53;               // staged raw block: the 53 aligned words that cover its 210 bytes
constexpr int RAW_STAGE_BYTES = BM * RAW_PITCH;              // one staging area (30 KB): refilled as soon as it has been converted
constexpr int Q_PITCH = 272, Q_SLOT_BYTES = BM * Q_PITCH;    // unpacked image of a block: 256 six-bit values as bytes | 16 scales; 17 x 16 B: conflict-free 16-byte reads
constexpr int EPI_WARPS = 16, PROD_WARPS = 8, PROD_THREADS = PROD_WARPS * 32, EPI_COLS = 16;
constexpr int MMA_WARP = EPI_WARPS + PROD_WARPS, LOAD_WARP = MMA_WARP + 2;   // two MMA issuers: one per column half
constexpr int THREADS = (LOAD_WARP + 1) * 32;                // 864
constexpr int SMEM_B = 0, SMEM_RAW = SMEM_B + B_SLOTS * TILE_BYTES, SMEM_Q = SMEM_RAW + RAW_STAGE_BYTES, SMEM_DX = SMEM_Q + Q_SLOTS * Q_SLOT_BYTES;
constexpr int SMEM_BARS = SMEM_DX + EPI_WARPS * 2 * 3 * 32 * 4, SMEM_TOTAL = SMEM_BARS + 256;   // DX: d_x and d, three buffers each per warp
constexpr int TMEM_COLS = 512;
constexpr uint32_t TM_HI = 0, TM_LO = 128, TM_A = 256, TM_A_STAGE = 64, TM_A_PIECE = 32;
constexpr uint32_t SBO_BYTES = 8 * KS * 2;                   // next 8-column group of core matrices (1024 B)
constexpr size_t FLAG_BYTES = 1024;
#ifndef Q6K_ISSUERS
#define Q6K_ISSUERS 2
#endif
#ifndef Q6K_WAIT_TFULL
#define Q6K_WAIT_TFULL 1
#endif
#ifndef Q6K_WAIT_TEMPTY
#define Q6K_WAIT_TEMPTY 1
#endif

struct Gemm6Params {
    const uint8_t *w;
    int64_t M, w_stride;
    int nb;
    const uint8_t *xq;      // fp16 activations, pre-tiled: [column tile][superblock][step][16 KB canonical K-major tile]
    const float *xd;        // [nb][N]
    float *dst;
    int64_t dst_stride;     // floats between dst columns
    int64_t N;
    int ksplit;
    int *flags;
    int32_t *dump_sumi;     // DUMP instantiation: [(n*M + m)*nb + b]
};

__device__ __forceinline__ uint32_t hfma2_u32(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// {s, s} as fp16 and {-1056 s, -1056 s}: the multiplier and the bias of one piece of a scale (|s| <= 15: all exact)
__device__ __forceinline__ void piece_consts(int s, uint32_t &mul, uint32_t &bias) {
    const __half h = __int2half_rn(s);
    const __half2 m = __half2half2(h), b = __hmul2(m, __half2half2(__ushort_as_half((unsigned short)0xe420)));   // -1056
    mul = *reinterpret_cast<const uint32_t *>(&m);
    bias = *reinterpret_cast<const uint32_t *>(&b);
}

// 16 weights of one 16-element sub-block: q = four words of 6-bit values (one per byte) -> 8 + 8 fp16x2 words (a*sh | a*sl)
__device__ __forceinline__ void unpack16(uint32_t *hi, uint32_t *lo, const uint32_t (&q)[4], int sc) {
    uint32_t mh, bh, ml, bl;
    piece_consts(sc >> 4, mh, bh);
    piece_consts(sc & 15, ml, bl);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t x01 = __byte_perm(q[i], 0x64646464u, 0x4140), x23 = __byte_perm(q[i], 0x64646464u, 0x4342);   // {1024 + q, 1024 + q'}
        hi[2 * i] = hfma2_u32(x01, mh, bh);      hi[2 * i + 1] = hfma2_u32(x23, mh, bh);
        lo[2 * i] = hfma2_u32(x01, ml, bl);      lo[2 * i + 1] = hfma2_u32(x23, ml, bl);
    }
}

// Stage the raw blocks of superblock `gkb` of the tile's 128 rows: warp `warp` (of 16) copies rows 8 warp .. 8 warp + 7 with
// 4-byte cp.async — the 53 aligned words that cover the row's 210-byte block, staged word j = global word j from
// (block address & ~3).  What is left of the misalignment is a per-row shift of 0 or 2 bytes (convert_rows).
// Row offsets are 32-bit (the launcher refuses matrices of 4 GB and more).
__device__ __forceinline__ void stage_raw(const Gemm6Params &p, uint8_t *stage, int warp, int lane, uint32_t row_off, int gkb) {
    // one row at a time, lanes = consecutive words
    uint32_t *dst0 = reinterpret_cast<uint32_t *>(stage + (8 * warp) * RAW_PITCH) + lane;
    const uint32_t last = (uint32_t)(p.M - 1) * (uint32_t)p.w_stride, step = (uint32_t)p.w_stride;
    uint32_t off = row_off;                                   // row 8 warp of the tile (warp-uniform), then one stride per row
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(p.w) + (uintptr_t)((off < last ? off : last) + (uint32_t)gkb * BLK_Q6K);
        const uint8_t *src = reinterpret_cast<const uint8_t *>(a & ~(uintptr_t)3) + 4 * lane;
        const uint32_t dst = smem_addr(dst0) + (uint32_t)i * RAW_PITCH;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
        if (lane + 32 < RAW_WORDS) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + 128u), "l"(src + 128) : "memory");
        off += step;
    }
}

// The same warp, once its copies have landed: staged raw rows -> the unpacked image the producers read.  A unit = sixteen l
// values of one 128-half (16 bytes each of ql[l], ql[l + 32], qh[l]) -> four 16-byte runs of 6-bit values, e = 0..3, at element
// 128 n + 32 e + l (dequantize_row_q6_K's order, L/ggml/src/ggml-quants.c:1762-1795).  Four units per row, one per thread:
// lane = row (8) + 8 x unit, so that the eight lanes of a quarter-warp read / write 16-byte pieces 15 (stage) and 17 (image)
// pieces apart — conflict-free.
__device__ __forceinline__ void convert_rows(const Gemm6Params &p, const uint8_t *stage, uint8_t *qimg, int warp, int lane, int64_t m0, int gkb) {
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
    const uint32_t w_lo = (uint32_t)reinterpret_cast<uintptr_t>(p.w) + (uint32_t)gkb * BLK_Q6K, stride_lo = (uint32_t)p.w_stride;
    auto shift_of = [&](int rl) {                                 // 0 or 16 bits: bit 1 of the row's block address
        int64_t row = m0 + rl;
        if (row >= p.M) row = p.M - 1;
        return ((w_lo + (uint32_t)row * stride_lo) & 2u) * 8u;
    };
    {
        const int rl = 8 * warp + (lane & 7), u = lane >> 3, n = u >> 1, jh = u & 1;
        const uint32_t sh = shift_of(rl);
        const uint8_t *A = stage + rl * RAW_PITCH;
        auto run16 = [&](int byte_off, uint32_t (&o)[4]) {       // 16 block bytes from staged offset byte_off (16-byte aligned) on
            const uint4 a = *reinterpret_cast<const uint4 *>(A + byte_off);
            const uint32_t b = *reinterpret_cast<const uint32_t *>(A + byte_off + 16);
            o[0] = __funnelshift_r(a.x, a.y, sh); o[1] = __funnelshift_r(a.y, a.z, sh);
            o[2] = __funnelshift_r(a.z, a.w, sh); o[3] = __funnelshift_r(a.w, b, sh);
        };
        uint32_t wa[4], wb[4], wh[4];
        run16(64 * n + 16 * jh, wa);
        run16(64 * n + 32 + 16 * jh, wb);
        run16(128 + 32 * n + 16 * jh, wh);
        uint4 e0, e1, e2, e3;
        uint32_t *o0 = &e0.x, *o1 = &e1.x, *o2 = &e2.x, *o3 = &e3.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            o0[i] = (wa[i] & 0x0f0f0f0fu) | ((wh[i] << 4) & 0x30303030u);
            o1[i] = (wb[i] & 0x0f0f0f0fu) | ((wh[i] << 2) & 0x30303030u);
            o2[i] = ((wa[i] >> 4) & 0x0f0f0f0fu) | (wh[i] & 0x30303030u);
            o3[i] = ((wb[i] >> 4) & 0x0f0f0f0fu) | ((wh[i] >> 2) & 0x30303030u);
        }
        uint8_t *out = qimg + rl * Q_PITCH + 128 * n + 16 * jh;
        *reinterpret_cast<uint4 *>(out) = e0;
        *reinterpret_cast<uint4 *>(out + 32) = e1;
        *reinterpret_cast<uint4 *>(out + 64) = e2;
        *reinterpret_cast<uint4 *>(out + 96) = e3;
    }
    {   // the 16 scales of each of the 8 rows: one word per lane
        const int rl = 8 * warp + (lane >> 2), c = lane & 3;
        const uint32_t *A = reinterpret_cast<const uint32_t *>(stage + rl * RAW_PITCH);
        reinterpret_cast<uint32_t *>(qimg + rl * Q_PITCH + 256)[c] = __funnelshift_r(A[48 + c], A[49 + c], shift_of(rl));
    }
    __syncwarp();                                                 // every lane has read the stage: it may be refilled
}

template <bool DUMP>
__global__ void __launch_bounds__(THREADS, 1)
gemm_q6k_tc_kernel(const __grid_constant__ Gemm6Params p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + SMEM_BARS);
    uint32_t *tmem_ptr_s = reinterpret_cast<uint32_t *>(bars + 24);
    const uint32_t bar0 = smem_addr(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };                  // [B_SLOTS]  producers + loader -> MMA
    auto done_bar = [&](int s) { return bar0 + 8u * (B_SLOTS + s); };      // [B_SLOTS]  MMA -> producers / loader
    auto tfull_bar = [&](int h) { return bar0 + 8u * (2 * B_SLOTS + h); };
    auto tempty_bar = [&](int h) { return bar0 + 8u * (2 * B_SLOTS + 2 + h); };
    auto qfull_bar = [&](int s) { return bar0 + 8u * (2 * B_SLOTS + 4 + s); };     // [Q_SLOTS] epilogue warps -> producers
    auto qempty_bar = [&](int s) { return bar0 + 8u * (2 * B_SLOTS + 4 + Q_SLOTS + s); };   // [Q_SLOTS] producers -> epilogue warps: the slot has been read

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t m0 = (int64_t)blockIdx.y * BM, n0 = (int64_t)blockIdx.x * BN;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int kb0 = (int)((int64_t)p.nb * blockIdx.z / p.ksplit);
    const int nb = (int)((int64_t)p.nb * (blockIdx.z + 1) / p.ksplit) - kb0, nsteps = SB_STEPS * nb;
    const int nh = n0 + HALF_N < p.N ? 2 : 1;               // column halves that hold real columns

    if (threadIdx.x == 0) {
        for (int s = 0; s < B_SLOTS; ++s) { mbar_init(full_bar(s), PROD_THREADS + 1); mbar_init(done_bar(s), Q6K_ISSUERS == 2 ? (uint32_t)nh : 1u); }
        for (int s = 0; s < Q_SLOTS; ++s) { mbar_init(qfull_bar(s), EPI_WARPS * 32); mbar_init(qempty_bar(s), PROD_THREADS); }
        for (int h = 0; h < 2; ++h) { mbar_init(tfull_bar(h), 1); mbar_init(tempty_bar(h), EPI_WARPS * 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(tmem_ptr_s)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_ptr_s;

    if (warp < EPI_WARPS) {
        // ===================== epilogue + raw weight feed =====================
        const int qd = warp & 3, g = warp >> 2;
        const int64_t row = m0 + 32 * qd + lane;
        // the two float scales of a (row, column, superblock) — d of the row's block, d_x of the column — are prefetched by the
        // warp itself with 4-byte cp.async into private buffers, two superblocks ahead (kb % 3), next to the raw blocks:
        // plain global loads in this loop are sunk to their use by the compiler and stalled every warp ~800 cycles per
        // superblock right where the tensor core waits for its accumulator back (profiles/r2_q6k_gemm_experiments.txt)
        float *dxw = reinterpret_cast<float *>(smem + SMEM_DX) + warp * 96;                              // [3][half 0: 16 | half 1: 16]
        uint32_t *dww = reinterpret_cast<uint32_t *>(smem + SMEM_DX + EPI_WARPS * 96 * 4) + warp * 96;   // [3][lane]: the word that holds d
        int64_t ncl = n0 + HALF_N * (lane >> 4) + EPI_COLS * g + (lane & 15);
        if (ncl >= p.N) ncl = p.N - 1;
        uint64_t acc[2][EPI_COLS / 2];
#pragma unroll
        for (int i = 0; i < EPI_COLS / 2; ++i) acc[0][i] = acc[1][i] = 0ull;
        const uintptr_t wd0 = reinterpret_cast<uintptr_t>(p.w) + (uintptr_t)((row < p.M ? row : p.M - 1) * p.w_stride) + (uintptr_t)kb0 * BLK_Q6K + 208;
        const float *xdk = p.xd + (int64_t)kb0 * p.N + ncl;
        auto prefetch_dw = [&](int kb) {
            const uintptr_t a = (wd0 + (uintptr_t)kb * BLK_Q6K) & ~(uintptr_t)3;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr(dww + (kb % 3) * 32 + lane)), "l"(a) : "memory");
        };
        auto prefetch_dx = [&](int kb) {
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr(dxw + (kb % 3) * 32 + lane)), "l"(xdk + (int64_t)kb * p.N) : "memory");
        };
        // the weights do not depend on the quantizer in front of this kernel: superblocks 0 and 1 are unpacked and superblock 2
        // is on its way before the wait
        uint8_t *stage = smem + SMEM_RAW;
        const int64_t srow = m0 + 8 * warp;                                   // first of the eight rows this warp stages
        const uint32_t row_off = (uint32_t)((srow < p.M ? srow : p.M - 1) * p.w_stride);   // (rows past M are clamped: valid memory, results never stored)
        stage_raw(p, stage, warp, lane, row_off, kb0);
        prefetch_dw(0);
        convert_rows(p, stage, smem + SMEM_Q, warp, lane, m0, kb0);
        mbar_arrive(qfull_bar(0));
        if (nb > 1) {
            stage_raw(p, stage, warp, lane, row_off, kb0 + 1);
            prefetch_dw(1);
            convert_rows(p, stage, smem + SMEM_Q + Q_SLOT_BYTES, warp, lane, m0, kb0 + 1);
            mbar_arrive(qfull_bar(1));
        }
        if (nb > 2) stage_raw(p, stage, warp, lane, row_off, kb0 + 2);
        asm volatile("griddepcontrol.wait;" ::: "memory");
        prefetch_dx(0);
        if (nb > 1) prefetch_dx(1);
        const uint32_t tbase = tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(EPI_COLS * g);
        const uint64_t k16 = pack2(16.0f, 16.0f);
        for (int kb = 0; kb < nb; ++kb) {
            // everything this thread has in flight (raw blocks of kb + 2, scales up to kb + 1) has landed.  Superblock kb + 2 is
            // unpacked into the slot of kb as soon as the producers have read that (they run up to four steps ahead of the tensor
            // core, so this is early in the MMAs of kb: the image is ready a whole superblock before it is needed and the warp is
            // back in time for this superblock's accumulators); then kb + 3 goes into the stage
            asm volatile("cp.async.wait_all;" ::: "memory");
            __syncwarp();
            if (kb + 2 < nb) {
                mbar_wait_long<Q4K_GEMM_WAIT>(qempty_bar(kb % Q_SLOTS), (uint32_t)((kb / Q_SLOTS) & 1));
                convert_rows(p, stage, smem + SMEM_Q + (kb % Q_SLOTS) * Q_SLOT_BYTES, warp, lane, m0, kb0 + kb + 2);
                mbar_arrive(qfull_bar(kb % Q_SLOTS));
                if (kb + 3 < nb) stage_raw(p, stage, warp, lane, row_off, kb0 + kb + 3);
                prefetch_dw(kb + 2);
                prefetch_dx(kb + 2);
            }
            const uint32_t dbits = dww[(kb % 3) * 32 + lane] >> (8u * (uint32_t)((wd0 + (uintptr_t)kb * BLK_Q6K) & 2u));
            const float dw = __half2float(__ushort_as_half((unsigned short)(dbits & 0xffffu)));
            const uint64_t d2 = pack2(dw, dw);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h >= nh) break;
                mbar_wait_long<Q6K_WAIT_TFULL>(tfull_bar(h), (uint32_t)(kb & 1));
                tc_fence_after();
                const uint64_t *dxk = reinterpret_cast<const uint64_t *>(dxw + (kb % 3) * 32 + 16 * h);
#pragma unroll
                for (int c8 = 0; c8 < EPI_COLS / 8; ++c8) {
                    uint32_t vh[8], vl[8];
                    const uint32_t ta = tbase + HALF_N * h + 8 * c8;
                    tmem_ld8(ta + TM_HI, vh); tmem_ld8(ta + TM_LO, vl);
                    tmem_ld_wait();
                    if (c8 == EPI_COLS / 8 - 1) {          // this half is in registers: hand the columns back
                        tc_fence_before();
                        mbar_arrive(tempty_bar(h));
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if constexpr (DUMP) {
                            const int64_t n = n0 + HALF_N * h + EPI_COLS * g + 8 * c8 + 2 * i;
                            if (row < p.M && n < p.N)
                                p.dump_sumi[(n * p.M + row) * p.nb + kb0 + kb] = 16 * (int)__uint_as_float(vh[2 * i]) + (int)__uint_as_float(vl[2 * i]);
                            if (row < p.M && n + 1 < p.N)
                                p.dump_sumi[((n + 1) * p.M + row) * p.nb + kb0 + kb] = 16 * (int)__uint_as_float(vh[2 * i + 1]) + (int)__uint_as_float(vl[2 * i + 1]);
                        }
                        // (float)sumi = RN(16 S_hi + S_lo);  t = d * sumi;  acc = fma(d_x, t, acc): pairs of columns
                        const uint64_t sf = fma2(pack2(__uint_as_float(vh[2 * i]), __uint_as_float(vh[2 * i + 1])), k16,
                                                 pack2(__uint_as_float(vl[2 * i]), __uint_as_float(vl[2 * i + 1])));
                        acc[h][4 * c8 + i] = fma2(dxk[4 * c8 + i], mul2(d2, sf), acc[h][4 * c8 + i]);
                    }
                }
            }
        }
        // split-K turnstile: parts add in K order (gemm_q4k_tc.cu)
        const int z = blockIdx.z;
        int *flag = p.ksplit > 1 ? p.flags + (blockIdx.y * gridDim.x + blockIdx.x) : nullptr;
        if (z > 0) {
            if (lane == 0) {
                int seen;
                do { asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(flag) : "memory"); } while (seen != z);
            }
            __syncwarp();
        }
        if (row < p.M) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h >= nh) break;
#pragma unroll
                for (int c = 0; c < EPI_COLS / 2; ++c) {
                    const int64_t n = n0 + HALF_N * h + EPI_COLS * g + 2 * c;
                    float a0, a1;
                    unpack2(acc[h][c], a0, a1);
                    if (z > 0) {
                        if (n < p.N) a0 = __fadd_rn(__ldcg(p.dst + n * p.dst_stride + row), a0);
                        if (n + 1 < p.N) a1 = __fadd_rn(__ldcg(p.dst + (n + 1) * p.dst_stride + row), a1);
                    }
                    if (n < p.N) p.dst[n * p.dst_stride + row] = a0;
                    if (n + 1 < p.N) p.dst[(n + 1) * p.dst_stride + row] = a1;
                }
            }
        }
        if (p.ksplit > 1 && z + 1 < p.ksplit) {
            __threadfence();
            asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");
            if (threadIdx.x == 0) asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(flag), "r"(z + 1) : "memory");
        }
    } else if (warp < EPI_WARPS + PROD_WARPS) {
        // ===================== producers: raw image -> A_hi / A_lo (fp16) in TMEM =====================
        const int t = threadIdx.x - EPI_WARPS * 32, pw = t >> 5;
        const int qd = pw & 3, k = pw >> 2, r = 32 * qd + lane;               // k: low (0) or high (1) nibbles / qh bit pairs
        const uint32_t ta_base = tmem + ((uint32_t)(32 * qd) << 16) + TM_A + 16u * k;
        uint4 scw = make_uint4(0, 0, 0, 0);
        for (int step = 0; step < nsteps; ++step) {
            const int kb = step >> 2, st = step & 3, n = st >> 1, hp = st & 1;
            if (st == 0) mbar_wait_long<Q4K_GEMM_WAIT>(qfull_bar(kb % Q_SLOTS), (uint32_t)((kb / Q_SLOTS) & 1));
            const uint8_t *qrow = smem + SMEM_Q + (kb % Q_SLOTS) * Q_SLOT_BYTES + r * Q_PITCH;
            if (st == 0) scw = *reinterpret_cast<const uint4 *>(qrow + 256);
            // e = 2k and 2k + 1, l = 16 hp .. 16 hp + 15: elements 128 n + 64 k + 16 hp .. and 32 further on
            const uint4 va = *reinterpret_cast<const uint4 *>(qrow + 128 * n + 64 * k + 16 * hp);
            const uint4 vb = *reinterpret_cast<const uint4 *>(qrow + 128 * n + 64 * k + 32 + 16 * hp);
            const uint32_t qa[4] = {va.x, va.y, va.z, va.w}, qb[4] = {vb.x, vb.y, vb.z, vb.w};
            // scales of sub-blocks 8n + 2e + hp: bytes 8n + 4k + hp and + 2 of the 16
            const uint32_t sw = n ? (k ? scw.w : scw.z) : (k ? scw.y : scw.x);
            const int sa = (int)(int8_t)(sw >> (8 * hp)), sb = (int)(int8_t)(sw >> (8 * hp + 16));
            uint32_t hi[16], lo[16];
            unpack16(hi, lo, qa, sa);
            unpack16(hi + 8, lo + 8, qb, sb);
            if (st == SB_STEPS - 1) mbar_arrive(qempty_bar(kb % Q_SLOTS));     // this thread has read the superblock's image
            // A stage step % 4 was last read by the MMAs of step - 4
            if (step >= A_STAGES) mbar_wait_long<Q4K_GEMM_WAIT>(done_bar((step - A_STAGES) % B_SLOTS), (uint32_t)(((step - A_STAGES) / B_SLOTS) & 1));
            tc_fence_after();
            const uint32_t ta = ta_base + (uint32_t)(step % A_STAGES) * TM_A_STAGE;
            tmem_st16(ta, hi);
            tmem_st16(ta + TM_A_PIECE, lo);
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(full_bar(step % B_SLOTS));
        }
    } else if (warp < LOAD_WARP) {
        // ===================== MMA issuers: warp MMA_WARP + h owns column half h =====================
        // (Q6K_ISSUERS = 1: one warp issues both halves, half 0 first — measured the same, profiles/r2_q6k_gemm_experiments.txt)
        const int hsel = warp - MMA_WARP;
        const int h_begin = Q6K_ISSUERS == 2 ? hsel : 0, h_end = Q6K_ISSUERS == 2 ? min(hsel + 1, nh) : (hsel == 0 ? nh : 0);
        if (h_begin < h_end) {
            const uint32_t idesc = umma_idesc(true);
            const uint32_t smem0 = smem_addr(smem);
            const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
            const uint64_t desc0 = umma_desc(0, SBO_BYTES) + (uint64_t)((smem0 + SMEM_B) >> 4);
            int bslot0 = 0;
            uint32_t bpar0 = 0u;
            for (int kb = 0; kb < nb; ++kb) {
                int bslot = bslot0;
                uint32_t bpar = bpar0;
                for (int h = h_begin; h < h_end; ++h) {
                    bslot = bslot0; bpar = bpar0;
                    const uint32_t d_hi = tm + TM_HI + h * HALF_N, d_lo = tm + TM_LO + h * HALF_N;
#pragma unroll
                    for (int st = 0; st < SB_STEPS; ++st) {
                        if (h == h_begin) mbar_wait_long<Q4K_GEMM_WAIT>(full_bar(bslot), bpar);
                        if (st == 0) mbar_wait_long<Q6K_WAIT_TEMPTY>(tempty_bar(h), (uint32_t)((kb & 1) ^ 1));
                        tc_fence_after();
                        const uint32_t ta = tm + TM_A + (uint32_t)st * TM_A_STAGE;                   // A stage = step % 4
                        const uint64_t bd0 = desc0 + (uint64_t)((uint32_t)bslot * (TILE_BYTES >> 4) + (uint32_t)h * ((HALF_N / 8) * SBO_BYTES >> 4));
                        if (elect_one()) {
#pragma unroll
                            for (int j = 0; j < KS / 16; ++j) {
                                const uint64_t b_desc = bd0 + (uint64_t)((j * 2 * LBO_BYTES) >> 4);
                                umma_f16_ts(d_hi, ta + 8 * j, b_desc, idesc, (uint32_t)((st | j) != 0));
                                umma_f16_ts(d_lo, ta + TM_A_PIECE + 8 * j, b_desc, idesc, (uint32_t)((st | j) != 0));
                            }
                            if (st == SB_STEPS - 1) umma_commit(tfull_bar(h));
                            if (Q6K_ISSUERS == 2 || h == nh - 1) umma_commit(done_bar(bslot));   // reads of the A stage and the B slot are done
                        }
                        __syncwarp();
                        if (++bslot == B_SLOTS) { bslot = 0; bpar ^= 1u; }
                    }
                }
                bslot0 = bslot; bpar0 = bpar;
            }
        }
    } else if (lane == 0) {
        // ===================== loader: the activation tiles, up to B_SLOTS steps ahead =====================
        const uint32_t smem0 = smem_addr(smem);
        asm volatile("griddepcontrol.wait;" ::: "memory");
        const uint8_t *xt = p.xq + (((int64_t)blockIdx.x * p.nb + kb0) << 16);
        for (int step = 0; step < nsteps; ++step) {
            if (step >= B_SLOTS) mbar_wait_long<Q4K_GEMM_WAIT>(done_bar(step % B_SLOTS), (uint32_t)(((step - B_SLOTS) / B_SLOTS) & 1));
            mbar_expect_tx(full_bar(step % B_SLOTS), TILE_BYTES);
            bulk_g2s(smem0 + SMEM_B + (step % B_SLOTS) * TILE_BYTES, xt + ((int64_t)step << 14), TILE_BYTES, full_bar(step % B_SLOTS));
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
    }
}

// F32 rows -> fp16 activations in the GEMM's tiled order + float d[nb][N]; the arithmetic of quantize_row_q8_K (shared device
// function, bit-exact), one warp per 256-block.  Element y of column c, superblock b lands in tile (c / 128, b), step
// 2 (y / 128) + (y / 16) % 2, slot s = 16 ((y / 32) % 4) + y % 16, at
//   ((c % 128) / 8) * 1024 + (s / 8) * 128 + (c % 8) * 16 + (s % 8) * 2        (canonical K-major core matrices, fp16)
__global__ void __launch_bounds__(256)
quantize_q8_f16_tiles_kernel(const float *__restrict__ x, uint8_t *__restrict__ qs, float *__restrict__ d, int64_t nb_per_row,
                             int64_t nblocks_total, int64_t x_row_stride_bytes, int64_t N, int *flags) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (flags && blockIdx.x == 0) flags[threadIdx.x] = 0;
    const int lane = threadIdx.x & 31;
    const int64_t blk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (blk >= nblocks_total) return;
    const int64_t col = blk / nb_per_row, b = blk - col * nb_per_row;
    const float *src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + col * x_row_stride_bytes) + b * QK_K;
    const float4 a = ld_act16(src + 4 * lane);
    const float4 c = ld_act16(src + 128 + 4 * lane);
    const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
    uint32_t w_lo, w_hi;
    int s_lo, s_hi;
    const float dq = quantize_block_q8_K<4>(v, lane, w_lo, w_hi, s_lo, s_hi);
    // int8 x 4 -> fp16 x 4: (q ^ 0x80) in the low byte of 0x64xx is 1024 + 128 + q; subtract 1152 (exact)
    const __half2 off = __half2half2(__ushort_as_half((unsigned short)0x6480));
    auto to_f16 = [&](uint32_t w) {
        const uint32_t u = w ^ 0x80808080u;
        const uint32_t p01 = __byte_perm(u, 0x64646464u, 0x4140), p23 = __byte_perm(u, 0x64646464u, 0x4342);
        const __half2 h01 = __hsub2(*reinterpret_cast<const __half2 *>(&p01), off), h23 = __hsub2(*reinterpret_cast<const __half2 *>(&p23), off);
        return make_uint2(*reinterpret_cast<const uint32_t *>(&h01), *reinterpret_cast<const uint32_t *>(&h23));
    };
    // lane l holds elements 4l .. 4l+3 (n = 0) and 128 + 4l .. (n = 1): e = l / 8, hp = (l / 4) % 2, j = 4 (l % 4)
    const int e = lane >> 3, hp = (lane >> 2) & 1, s0 = 16 * e + 4 * (lane & 3);
    const int64_t ntile = col >> 7, nl = col & 127;
    uint8_t *tile = qs + ((ntile * nb_per_row + b) << 16) + (nl >> 3) * 1024 + (s0 >> 3) * 128 + (nl & 7) * 16 + (s0 & 7) * 2;
    *reinterpret_cast<uint2 *>(tile + hp * TILE_BYTES) = to_f16(w_lo);
    *reinterpret_cast<uint2 *>(tile + (2 + hp) * TILE_BYTES) = to_f16(w_hi);
    if (lane == 0) d[b * N + col] = dq;
}

int choose_ksplit6(int64_t tiles, int64_t nb) {              // as gemm_q4k_tc.cu: split only long K while every CTA stays resident
    if (knobs().gemm_ksplit >= 1) return (int)std::min<int64_t>(knobs().gemm_ksplit, nb);
    const int sms = sm_count();
    int best = 1;
    double best_cost = (double)nb;
    for (int s = 2; s <= 4; ++s) {
        if (tiles * s > sms || nb / s < 4) break;
        const double cost = (double)((nb + s - 1) / s) + 5.0 * (s - 1);       // a superblock costs ~2.5 us here, a hand-off ~12
        if (cost < best_cost - 2.0) { best_cost = cost; best = s; }
    }
    return best;
}

}  // namespace

// workspace of the Q6_K prefill path: fp16 tiles | float d | split-K turnstiles
size_t gemm_q6k_workspace_bytes(int64_t K, int64_t N) {
    const size_t nb = (size_t)(K / QK_K), ntiles = ((size_t)N + BN - 1) / BN;
    return ntiles * BN * (size_t)K * 2 + ((nb * (size_t)N * 4 + 255) & ~(size_t)255) + FLAG_BYTES;
}

// F32 activations: quantize into `workspace` (gemm_q6k_workspace_bytes), then the tensor-core GEMM.  Two launches.
cudaError_t launch_gemm_q6k_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x, int64_t N,
                                int64_t x_row_stride_bytes, float *dst, int64_t dst_col_stride_bytes, void *workspace,
                                cudaStream_t stream, int32_t *dump_sumi) {
    if (N == 0 || M == 0) return cudaSuccess;
    const int64_t nb = K / QK_K;
    const int64_t ntiles = (N + BN - 1) / BN, mtiles = (M + BM - 1) / BM;
    if (mtiles > 65535 || M * w_row_stride_bytes >= ((int64_t)1 << 32)) return cudaErrorNotSupported;     // (the caller runs the decode kernel instead)
    uint8_t *qs = reinterpret_cast<uint8_t *>(workspace);
    float *d = reinterpret_cast<float *>(qs + ntiles * BN * K * 2);
    int *flags = reinterpret_cast<int *>(reinterpret_cast<char *>(d) + ((nb * (size_t)N * 4 + 255) & ~(size_t)255));
    const int ksplit = choose_ksplit6(ntiles * mtiles, nb);
    const int64_t total = nb * N;
    cudaLaunchAttribute pdl_attr[1];
    pdl_attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pdl_attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaError_t e;
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)((total + 7) / 8));
        cfg.blockDim = dim3(256);
        cfg.stream = stream;
        cfg.attrs = pdl_attr;
        cfg.numAttrs = knobs().pdl ? 1 : 0;
        int *qflags = ksplit > 1 ? flags : nullptr;
        e = cudaLaunchKernelEx(&cfg, quantize_q8_f16_tiles_kernel, x, qs, d, nb, total, x_row_stride_bytes, N, qflags);
        count_launch();
        if (e != cudaSuccess) return e;
    }
    Gemm6Params p;
    p.w = (const uint8_t *)w; p.M = M; p.w_stride = w_row_stride_bytes; p.nb = (int)nb;
    p.xq = qs; p.xd = d; p.dst = dst; p.dst_stride = dst_col_stride_bytes / 4; p.N = N;
    p.ksplit = ksplit; p.flags = flags; p.dump_sumi = dump_sumi;
    auto kern = dump_sumi ? gemm_q6k_tc_kernel<true> : gemm_q6k_tc_kernel<false>;
    if ((e = ensure_max_dynamic_smem((const void *)kern, SMEM_TOTAL)) != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ntiles, (unsigned)mtiles, (unsigned)ksplit);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = SMEM_TOTAL;
    cfg.stream = stream;
    cfg.attrs = pdl_attr;
    cfg.numAttrs = knobs().pdl ? 1 : 0;
    e = cudaLaunchKernelEx(&cfg, kern, p);
    count_launch();
    return e;
}

}  // namespace q4k
