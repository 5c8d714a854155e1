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
// CTA = 128 weight rows x 128 activation columns, 26 warps, warp-specialised; a pipeline step is 64 elements of K (4 per
// superblock):
//   warps 0-15   epilogue (thread <-> TMEM lane = weight row, column group g = warp / 4 of both 64-column halves) AND the raw
//                weight feed: Q6_K rows are 210-byte blocks at 2-byte alignment, which neither TMA nor cp.async can copy, so
//                these warps — idle most of a superblock — load them as aligned words, funnel-shift them into place and store
//                a 16-byte aligned image (row pitch 240 B: conflict-free 16-byte reads) two superblocks ahead.
//   warps 16-23  producers: thread <-> weight row (quadrant = warp % 4) and K half (warp / 4) of the step: 6-bit unpack,
//                two HFMA2 per pair, tcgen05.st into one of four A stages.
//   warp 24      allocates TMEM; one lane issues the MMAs (M = 128, N = 64, K = 16) and commits to mbarriers.
//   warp 25      one lane: bulk copies of the activation tiles (pre-tiled by the quantizer in exactly the shared-memory order).
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
constexpr int B_SLOTS = 6, A_STAGES = 4, RAW_SLOTS = 3;
constexpr int TILE_BYTES = BN * KS * 2;                      // 16 KB: the fp16 activation tile of a step
constexpr int RAW_PITCH = 240, RAW_WORDS = 53;               // realigned block image: 53 words (212 B), row pitch 240 B
constexpr int RAW_SLOT_BYTES = BM * RAW_PITCH;
constexpr int EPI_WARPS = 16, PROD_WARPS = 8, PROD_THREADS = PROD_WARPS * 32, EPI_COLS = 16;
constexpr int MMA_WARP = EPI_WARPS + PROD_WARPS, LOAD_WARP = MMA_WARP + 1;
constexpr int THREADS = (LOAD_WARP + 1) * 32;                // 832
constexpr int SMEM_B = 0, SMEM_RAW = SMEM_B + B_SLOTS * TILE_BYTES, SMEM_DX = SMEM_RAW + RAW_SLOTS * RAW_SLOT_BYTES;
constexpr int SMEM_BARS = SMEM_DX + EPI_WARPS * 2 * 32 * 4, SMEM_TOTAL = SMEM_BARS + 256;
constexpr int TMEM_COLS = 512;
constexpr uint32_t TM_HI = 0, TM_LO = 128, TM_A = 256, TM_A_STAGE = 64, TM_A_PIECE = 32;
constexpr uint32_t SBO_BYTES = 8 * KS * 2;                   // next 8-column group of core matrices (1024 B)
constexpr size_t FLAG_BYTES = 1024;

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

__device__ __forceinline__ uint32_t ldg_nc_u32(const void *p) {
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

// the raw image of superblock `gkb` of the tile's 128 rows: warp `warp` (of 16) realigns rows 8 warp .. 8 warp + 7
__device__ __forceinline__ void feed_raw(const Gemm6Params &p, uint8_t *slot, int warp, int lane, int64_t m0, int gkb) {
    uint32_t a0[8], a1[8];
    uint32_t sh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int64_t row = m0 + 8 * warp + i;
        if (row >= p.M) row = p.M - 1;                       // rows past M: valid memory, results never stored
        const uintptr_t base = reinterpret_cast<uintptr_t>(p.w) + (uintptr_t)(row * p.w_stride) + (uintptr_t)gkb * BLK_Q6K;
        const uint8_t *al = reinterpret_cast<const uint8_t *>(base & ~(uintptr_t)3);
        sh[i] = (uint32_t)(base & 2u) * 8u;
        a0[i] = ldg_nc_u32(al + 4 * lane);
        a1[i] = lane + 32 < RAW_WORDS ? ldg_nc_u32(al + 4 * (lane + 32)) : 0u;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        // word k of the image = bytes [4k, 4k + 4) of the block = funnel of aligned words k, k + 1 (the word past the block's
        // last one is never loaded: lanes >= 21 of the second pass hold zero)
        uint32_t n0 = __shfl_down_sync(0xffffffffu, a0[i], 1);
        const uint32_t first1 = __shfl_sync(0xffffffffu, a1[i], 0), n1 = __shfl_down_sync(0xffffffffu, a1[i], 1);
        if (lane == 31) n0 = first1;
        uint32_t *dst = reinterpret_cast<uint32_t *>(slot + (8 * warp + i) * RAW_PITCH);
        dst[lane] = __funnelshift_r(a0[i], n0, sh[i]);
        if (lane + 32 < RAW_WORDS) dst[lane + 32] = __funnelshift_r(a1[i], lane + 33 < RAW_WORDS ? n1 : 0u, sh[i]);
    }
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
    auto rawfull_bar = [&](int s) { return bar0 + 8u * (2 * B_SLOTS + 4 + s); };   // [RAW_SLOTS] epilogue warps -> producers

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t m0 = (int64_t)blockIdx.y * BM, n0 = (int64_t)blockIdx.x * BN;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int kb0 = (int)((int64_t)p.nb * blockIdx.z / p.ksplit);
    const int nb = (int)((int64_t)p.nb * (blockIdx.z + 1) / p.ksplit) - kb0, nsteps = SB_STEPS * nb;
    const int nh = n0 + HALF_N < p.N ? 2 : 1;               // column halves that hold real columns

    if (threadIdx.x == 0) {
        for (int s = 0; s < B_SLOTS; ++s) { mbar_init(full_bar(s), PROD_THREADS + 1); mbar_init(done_bar(s), 1); }
        for (int s = 0; s < RAW_SLOTS; ++s) mbar_init(rawfull_bar(s), EPI_WARPS * 32);
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
        float *dxw = reinterpret_cast<float *>(smem + SMEM_DX) + warp * 64;   // [buffer][half 0: 16 | half 1: 16]
        int64_t ncl = n0 + HALF_N * (lane >> 4) + EPI_COLS * g + (lane & 15);
        if (ncl >= p.N) ncl = p.N - 1;
        uint64_t acc[2][EPI_COLS / 2];
#pragma unroll
        for (int i = 0; i < EPI_COLS / 2; ++i) acc[0][i] = acc[1][i] = 0ull;
        // the weights do not depend on the quantizer in front of this kernel: the first two superblocks go out before the wait
        for (int kb = 0; kb < min(nb, RAW_SLOTS - 1); ++kb) {
            feed_raw(p, smem + SMEM_RAW + kb * RAW_SLOT_BYTES, warp, lane, m0, kb0 + kb);
            mbar_arrive(rawfull_bar(kb));
        }
        // d of this thread's row, one superblock ahead (2-byte loads at block offset 208; not from the raw image, whose slots
        // the faster warps recycle while a slower one is still on the previous superblock)
        const uint8_t *wd = p.w + (row < p.M ? row : p.M - 1) * p.w_stride + (int64_t)kb0 * BLK_Q6K + 208;
        unsigned short dh_next = __ldg(reinterpret_cast<const unsigned short *>(wd));
        const float *xdk = p.xd + (int64_t)kb0 * p.N;
        asm volatile("griddepcontrol.wait;" ::: "memory");
        dxw[lane] = xdk[ncl];
        __syncwarp();
        const uint32_t tbase = tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(EPI_COLS * g);
        const uint64_t k16 = pack2(16.0f, 16.0f);
        for (int kb = 0; kb < nb; ++kb) {
            // superblock kb + 2 takes the raw slot of kb - 1, whose last MMA completed before the tfull this warp saw last
            if (kb + RAW_SLOTS - 1 < nb) {
                const int kn = kb + RAW_SLOTS - 1;
                feed_raw(p, smem + SMEM_RAW + (kn % RAW_SLOTS) * RAW_SLOT_BYTES, warp, lane, m0, kb0 + kn);
                mbar_arrive(rawfull_bar(kn % RAW_SLOTS));
            }
            float dx_next = 0.0f;
            const float dw = __half2float(__ushort_as_half(dh_next));
            const uint64_t d2 = pack2(dw, dw);
            if (kb + 1 < nb) {
                dx_next = xdk[(int64_t)(kb + 1) * p.N + ncl];
                dh_next = __ldg(reinterpret_cast<const unsigned short *>(wd + (int64_t)(kb + 1) * BLK_Q6K));
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h >= nh) break;
                mbar_wait_long<Q4K_GEMM_WAIT_TFULL>(tfull_bar(h), (uint32_t)(kb & 1));
                tc_fence_after();
                const uint64_t *dxk = reinterpret_cast<const uint64_t *>(dxw + (kb & 1) * 32 + 16 * h);
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
            dxw[((kb + 1) & 1) * 32 + lane] = dx_next;
            __syncwarp();
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
            if (st == 0) mbar_wait_long<Q4K_GEMM_WAIT>(rawfull_bar(kb % RAW_SLOTS), (uint32_t)((kb / RAW_SLOTS) & 1));
            const uint8_t *raw = smem + SMEM_RAW + (kb % RAW_SLOTS) * RAW_SLOT_BYTES + r * RAW_PITCH;
            if (st == 0) scw = *reinterpret_cast<const uint4 *>(raw + 192);
            const uint4 la = *reinterpret_cast<const uint4 *>(raw + 64 * n + 16 * hp);            // ql[l],      l = 16 hp ..
            const uint4 lb = *reinterpret_cast<const uint4 *>(raw + 64 * n + 32 + 16 * hp);       // ql[32 + l]
            const uint4 hq = *reinterpret_cast<const uint4 *>(raw + 128 + 32 * n + 16 * hp);      // qh[l]
            const uint32_t wa[4] = {la.x, la.y, la.z, la.w}, wb[4] = {lb.x, lb.y, lb.z, lb.w}, wh[4] = {hq.x, hq.y, hq.z, hq.w};
            uint32_t qa[4], qb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {        // e = 2k: nibble k of ql[l], qh bits 4k, 4k+1;  e = 2k + 1: nibble k of ql[32 + l], bits 4k+2, 4k+3
                const uint32_t h4 = k ? wh[i] >> 4 : wh[i];
                qa[i] = ((k ? wa[i] >> 4 : wa[i]) & 0x0f0f0f0fu) | ((h4 << 4) & 0x30303030u);
                qb[i] = ((k ? wb[i] >> 4 : wb[i]) & 0x0f0f0f0fu) | ((h4 << 2) & 0x30303030u);
            }
            // scales of sub-blocks 8n + 2e + hp: bytes 8n + 4k + hp and + 2 of the 16
            const uint32_t sw = n ? (k ? scw.w : scw.z) : (k ? scw.y : scw.x);
            const int sa = (int)(int8_t)(sw >> (8 * hp)), sb = (int)(int8_t)(sw >> (8 * hp + 16));
            uint32_t hi[16], lo[16];
            unpack16(hi, lo, qa, sa);
            unpack16(hi + 8, lo + 8, qb, sb);
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
    } else if (warp == MMA_WARP) {
        // ===================== MMA issuer =====================
        const uint32_t idesc = umma_idesc(true);
        const uint32_t smem0 = smem_addr(smem);
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
        const uint64_t desc0 = umma_desc(0, SBO_BYTES);
        // per superblock: the 32 MMAs of column half 0 (four K steps x two pieces x four K = 16 slices), then those of half 1;
        // the epilogue drains one half while the tensor core works on the other
        for (int kb = 0; kb < nb; ++kb) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h >= nh) break;
#pragma unroll
                for (int st = 0; st < SB_STEPS; ++st) {
                    const int step = SB_STEPS * kb + st, bslot = step % B_SLOTS;
                    if (h == 0) mbar_wait_long<Q4K_GEMM_WAIT>(full_bar(bslot), (uint32_t)((step / B_SLOTS) & 1));
                    if (st == 0) mbar_wait_long<Q4K_GEMM_WAIT_TEMPTY>(tempty_bar(h), (uint32_t)((kb & 1) ^ 1));
                    tc_fence_after();
                    const uint32_t ta = tm + TM_A + (uint32_t)(step % A_STAGES) * TM_A_STAGE;
                    const uint64_t bd0 = desc0 + (uint64_t)((smem0 + SMEM_B + (uint32_t)bslot * TILE_BYTES) >> 4);
                    if (elect_one()) {
#pragma unroll
                        for (int o = 0; o < 2; ++o) {
#pragma unroll
                            for (int j = 0; j < KS / 16; ++j) {
                                const uint64_t b_desc = bd0 + (uint64_t)((h * (HALF_N / 8) * SBO_BYTES + j * 2 * LBO_BYTES) >> 4);
                                umma_f16_ts(tm + (o ? TM_LO : TM_HI) + h * HALF_N, ta + o * TM_A_PIECE + 8 * j, b_desc, idesc, (uint32_t)((st | j) != 0));
                            }
                        }
                        if (st == SB_STEPS - 1) umma_commit(tfull_bar(h));
                        if (h == nh - 1) umma_commit(done_bar(bslot));
                    }
                    __syncwarp();
                }
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
    const float4 a = reinterpret_cast<const float4 *>(src)[lane];
    const float4 c = reinterpret_cast<const float4 *>(src)[32 + lane];
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
        const double cost = (double)((nb + s - 1) / s) + 10.0 * (s - 1);
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
    if (mtiles > 65535) return cudaErrorNotSupported;
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
