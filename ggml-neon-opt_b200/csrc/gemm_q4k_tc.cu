// gemm_q4k_tc.cu — prefill GEMM (batch > 8) on the 5th-generation tensor cores: tcgen05.mma kind::i8, TMEM accumulators.
// dst[n][m] = sum_blk  d_x[n,blk] * ( d_w[m,blk] * sumi[m,n,blk]  -  dmin_w[m,blk] * summ[m,n,blk] )
// which is ggml_compute_forward_mul_mat for Q4_K x F32 at ne11 = N (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401) with
// the integer stage of ggml_vec_dot_q4_K_q8_K (L/ggml/src/ggml-cpu/quants.c:575-611) kept EXACT:
//   sumi = sum_j sc[j] * p[j]        p[j] = sum_{l<32} q4[32j+l] * q8[32j+l]
//   summ = sum_j m[j]  * (sum of the 32 q8 of sub-block j)
//
// How the per-sub-block 6-bit scales get through an int8 tensor core exactly.  sc <= 63 and q4 <= 15, so
// sc*q4 needs 10 bits.  Split the SCALE, not the product:  sc = 8*sh + sl  (sh, sl <= 7).  Then sh*q4 and
// sl*q4 are <= 105 and fit a signed byte, a packed multiply of four nibble-bytes by sh never carries between
// bytes, and   sumi = 8 * (A_hi . q8) + (A_lo . q8)   with A_hi = sh*q4, A_lo = sl*q4 accumulated over the
// whole 256-wide superblock in two TMEM accumulators.  The mins ride the tensor core too: summ = A_m . q8 with
// A_m[k] = m[sub-block of k] (<= 63), a third accumulator.  One TMEM drain per superblock instead of one per
// 32-wide sub-block; the fp32 scales d_w*d_x, dmin_w*d_x are applied by the epilogue warps per superblock,
// sequentially over superblocks (the NEON order of the reference; within 1e-5 of the AVX2 order).
//
// CTA = 128 weight rows x 128 activation columns, 21 warps, warp-specialised:
//   warps 0-15  epilogue: thread <-> TMEM lane (= weight row); group g = warp / 4 owns columns 32g..32g+31 (groups
//               0,1 = accumulator half 0, groups 2,3 = half 1).  tcgen05.ld 8 columns x 3 accumulators at a time,
//               double-buffered, 32 fp32 results per thread in registers.
//   warps 16-19 producers: coalesced 16-byte cp.async of (a) whole block_q4_K rows of the tile's superblock into a raw
//               ring and (b) the int8 activation tile, which is stored PRE-TILED in global memory in exactly the
//               shared-memory order, straight into its B slot; then thread <-> weight row: unpack scales, write the
//               three int8 A tiles of half a superblock (K = 128) in the canonical K-major no-swizzle UMMA layout
//               (8-row x 16-byte core matrices).
//   warp 20     allocates TMEM; one lane issues the MMAs (M = 128, N = 64, K = 32) and commits to mbarriers.
// Pipelines: A stages (2 x 48 KB), B slots (4 x 16 KB), raw ring (3 x 18 KB), 3 steps of loads in flight; the two
// column halves of the accumulators alternate between the tensor core and their epilogue groups (the MMAs of half 1
// overlap the drain of half 0).
//
// Activations (quantize_q8_planar_kernel below): the same int8 / float d that quantize_row_q8_K produces (shared
// device function, bit-exact), int8 in tile order + float d[nb][N].  bsums are not needed: the mins use the MMA.
#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

constexpr int BM = 128, BN = 128, BK = 128;          // CTA tile; BK = K bytes per pipeline stage (half a superblock)
constexpr int HALF_N = 64;                            // columns per accumulator half / per MMA
constexpr int A_STAGES = 2;                           // unpacked A tiles (A_hi, A_lo, A_m): producers -> MMA
constexpr int B_SLOTS = 4;                            // activation tiles, written in place by cp.async
constexpr int RAW_SLOTS = 3;                          // raw block_q4_K bytes of 128 rows x one superblock (coalesced cp.async)
constexpr int PREFETCH = 3;                           // pipeline steps of global loads in flight per producer thread
constexpr int TILE_BYTES = BM * BK;                   // 16 KB: one int8 operand tile of a stage
constexpr int A_STAGE_BYTES = 3 * TILE_BYTES;         // A_hi, A_lo, A_m
constexpr int RAW_ROW_BYTES = BLK_Q4K, RAW_SLOT_BYTES = BM * RAW_ROW_BYTES;
constexpr int SMEM_A = 0, SMEM_B = SMEM_A + A_STAGES * A_STAGE_BYTES, SMEM_RAW = SMEM_B + B_SLOTS * TILE_BYTES;
constexpr int SMEM_DX = SMEM_RAW + RAW_SLOTS * RAW_SLOT_BYTES, SMEM_BARS = SMEM_DX + 1024;
constexpr int SMEM_TOTAL = SMEM_BARS + 256;
constexpr int EPI_WARPS = 16, PROD_WARPS = 4;        // 4 epilogue groups of 4 warps (32 columns each), 2 groups per accumulator half
constexpr int EPI_COLS = BN / (EPI_WARPS / 4);        // 32 columns per epilogue thread
constexpr int GEMM_THREADS = (EPI_WARPS + PROD_WARPS + 1) * 32;   // 672
constexpr int TMEM_COLS = 512;                        // 3 accumulators x 128 columns, rounded to a power of two
constexpr uint32_t SBO_BYTES = 8 * BK;                // 1024: next 8-row group of core matrices
constexpr uint32_t LBO_BYTES = 128;                   // next 16-byte K chunk (core matrices of a row group are contiguous)

struct GemmParams {
    const uint8_t *w;
    int64_t M, w_stride;
    int nb;
    const int8_t *xq;       // int8 activations, pre-tiled: [column tile][superblock][half][16 KB canonical B tile]
    const float *xd;        // [nb][Nld]
    int64_t N, Nld;
    float *dst;
    int64_t dst_stride;     // floats between dst columns
};

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// shared-memory matrix descriptor, K-major, SWIZZLE_NONE ("interleave"): 8-row x 16-byte core matrices,
// LBO = byte distance between the two 16-byte K chunks of one MMA, SBO = byte distance between 8-row groups
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffffu) >> 4);                    // start address, bits [0,14)
    d |= (uint64_t)(LBO_BYTES >> 4) << 16;                       // leading byte offset, bits [16,30)
    d |= (uint64_t)(SBO_BYTES >> 4) << 32;                       // stride byte offset, bits [32,46)
    d |= (uint64_t)1 << 46;                                      // descriptor version 1 (sm_100)
    return d;                                                    // base offset 0, layout type 0 = no swizzle
}

// instruction descriptor for kind::i8: D = S32, A = B = signed 8-bit, both K-major, M = 128, N = 64
__device__ __forceinline__ uint32_t umma_idesc_i8() {
    uint32_t d = 0;
    d |= 2u << 4;                   // c_format: S32
    d |= 1u << 7;                   // a_format: signed 8-bit
    d |= 1u << 10;                  // b_format: signed 8-bit
    d |= (uint32_t)(HALF_N >> 3) << 17;
    d |= (uint32_t)(BM >> 4) << 24;
    return d;
}

__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on the mbarrier when every MMA issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

}  // namespace

namespace {

__device__ __forceinline__ uint4 ldg16(const void *p) { return *reinterpret_cast<const uint4 *>(p); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Global -> shared copies of one pipeline step, 16 bytes per cp.async, consecutive threads on consecutive chunks
// (coalesced; the destination is linear in the chunk index for both operands):
//   B tile of `step`      : 1024 chunks, the activations are stored pre-tiled in exactly the shared-memory order
//   raw weights of `sb`   : 128 rows x 9 chunks (whole block_q4_K), issued together with the B tile of step 2*sb
__device__ __forceinline__ void producer_issue(const GemmParams &p, uint32_t smem0, int t, int64_t m0, int64_t ntile, int step) {
    const int kb = step >> 1, hh = step & 1;
    const int8_t *bsrc = p.xq + (((ntile * p.nb + kb) * 2 + hh) << 14) + 16 * t;
    const uint32_t bdst = smem0 + SMEM_B + (step % B_SLOTS) * TILE_BYTES + 16 * t;
#pragma unroll
    for (int i = 0; i < 8; ++i) cp_async16(bdst + i * 2048, bsrc + i * 2048);
    if (hh == 0) {
        const uint32_t raw = smem0 + SMEM_RAW + (kb % RAW_SLOTS) * RAW_SLOT_BYTES;
#pragma unroll
        for (int i = 0; i < 9; ++i) {
            const int c = i * 128 + t, r = c / 9, ch = c - 9 * r;
            int64_t row = m0 + r;
            if (row >= p.M) row = p.M - 1;
            cp_async16(raw + 16 * c, p.w + row * p.w_stride + (int64_t)kb * BLK_Q4K + 16 * ch);
        }
    }
}

// the thread's raw row (in shared memory) -> three int8 A tiles of the stage
__device__ __forceinline__ void producer_unpack(uint8_t *a_stage, const uint8_t *raw, int t, int hh) {
    const uint4 hdr = *reinterpret_cast<const uint4 *>(raw);
    uint4 q[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] = *reinterpret_cast<const uint4 *>(raw + 16 + 64 * hh + 16 * i);
    const uint4 sm = unpack_scales_mins(hdr);
    const uint32_t scw = hh ? sm.y : sm.x, mw = hh ? sm.w : sm.z;
    uint8_t *row_base = a_stage + (t >> 3) * SBO_BYTES + (t & 7) * 16;
#pragma unroll
    for (int gi = 0; gi < 2; ++gi) {
        const uint32_t qw[8] = {q[2 * gi].x, q[2 * gi].y, q[2 * gi].z, q[2 * gi].w,
                                q[2 * gi + 1].x, q[2 * gi + 1].y, q[2 * gi + 1].z, q[2 * gi + 1].w};
#pragma unroll
        for (int nib = 0; nib < 2; ++nib) {             // 0: low nibbles = sub-block 2g, 1: high nibbles = sub-block 2g + 1
            const uint32_t sc = (scw >> (16 * gi + 8 * nib)) & 0xffu, mn = (mw >> (16 * gi + 8 * nib)) & 0xffu;
            const uint32_t sh = sc >> 3, sl = sc & 7u, mrep = mn * 0x01010101u;
            uint32_t v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = (nib ? (qw[i] >> 4) : qw[i]) & 0x0f0f0f0fu;
#pragma unroll
            for (int c = 0; c < 2; ++c) {               // two 16-element K chunks per sub-block
                uint8_t *dst = row_base + (4 * gi + 2 * nib + c) * LBO_BYTES;
                *reinterpret_cast<uint4 *>(dst) = make_uint4(v[4 * c] * sh, v[4 * c + 1] * sh, v[4 * c + 2] * sh, v[4 * c + 3] * sh);
                *reinterpret_cast<uint4 *>(dst + TILE_BYTES) = make_uint4(v[4 * c] * sl, v[4 * c + 1] * sl, v[4 * c + 2] * sl, v[4 * c + 3] * sl);
                *reinterpret_cast<uint4 *>(dst + 2 * TILE_BYTES) = make_uint4(mrep, mrep, mrep, mrep);
            }
        }
    }
}

__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_q4k_tc_kernel(const GemmParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    float *dx_s = reinterpret_cast<float *>(smem + SMEM_DX);            // [half][buffer][64]
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + SMEM_BARS);
    uint32_t *tmem_ptr_s = reinterpret_cast<uint32_t *>(bars + 24);
    const uint32_t bar0 = smem_addr(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };                  // [B_SLOTS]  producers -> MMA, one per pipeline step slot
    auto emptya_bar = [&](int s) { return bar0 + 8u * (4 + s); };          // [A_STAGES] MMA -> producers
    auto emptyb_bar = [&](int s) { return bar0 + 8u * (6 + s); };          // [B_SLOTS]  MMA -> producers
    auto tfull_bar = [&](int h) { return bar0 + 8u * (10 + h); };
    auto tempty_bar = [&](int h) { return bar0 + 8u * (12 + h); };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t m0 = (int64_t)blockIdx.y * BM, n0 = (int64_t)blockIdx.x * BN;
    const int nb = p.nb, nsteps = 2 * nb;
    const int64_t K = (int64_t)nb * QK_K;

    if (threadIdx.x == 0) {
        for (int s = 0; s < B_SLOTS; ++s) { mbar_init(full_bar(s), PROD_WARPS * 32); mbar_init(emptyb_bar(s), 1); }
        for (int s = 0; s < A_STAGES; ++s) mbar_init(emptya_bar(s), 1);
        for (int h = 0; h < 2; ++h) { mbar_init(tfull_bar(h), 1); mbar_init(tempty_bar(h), (EPI_WARPS / 2) * 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == EPI_WARPS + PROD_WARPS) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(tmem_ptr_s)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_ptr_s;

    if (warp < EPI_WARPS) {
        // ===================== epilogue: TMEM -> registers, per-superblock fp32 scales =====================
        const int qd = warp & 3, grp = warp >> 2, h = grp >> 1;               // lane quadrant, column group (32 wide), accumulator half
        const int th = threadIdx.x & 255;                                      // thread within the half (two groups)
        const int64_t row = m0 + 32 * qd + lane;
        const uint8_t *wrow = p.w + (row < p.M ? row : p.M - 1) * p.w_stride;
        float *dx = dx_s + h * 128;
        int64_t ncl = n0 + HALF_N * h + (th & 63);
        if (ncl >= p.N) ncl = p.N - 1;
        float acc[EPI_COLS];
#pragma unroll
        for (int i = 0; i < EPI_COLS; ++i) acc[i] = 0.0f;
        uint32_t dd_next = *reinterpret_cast<const uint32_t *>(wrow);
        if (th < 64) dx[th] = p.xd[ncl];                                       // superblock 0 -> buffer 0
        const uint32_t tbase = tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(EPI_COLS * grp);
        const int cofs = EPI_COLS * (grp & 1);                                 // this group's columns inside the half
        for (int kb = 0; kb < nb; ++kb) {
            const uint32_t dd = dd_next;
            if (kb + 1 < nb) dd_next = *reinterpret_cast<const uint32_t *>(wrow + (int64_t)(kb + 1) * BLK_Q4K);
            mbar_wait(tfull_bar(h), (uint32_t)(kb & 1));
            tc_fence_after();
            asm volatile("bar.sync %0, 256;" ::"r"(1 + h) : "memory");        // everyone of this half is done with superblock kb - 1
            if (th < 64 && kb + 1 < nb) dx[((kb + 1) & 1) * 64 + th] = p.xd[(int64_t)(kb + 1) * p.Nld + ncl];
            const float2 dw = __half22float2(*reinterpret_cast<const __half2 *>(&dd));   // {d, dmin} of this row's superblock
            const float *dxk = dx + (kb & 1) * 64 + cofs;
            uint32_t vh[2][8], vl[2][8], vm[2][8];
            tmem_ld8(tbase, vh[0]); tmem_ld8(tbase + 128, vl[0]); tmem_ld8(tbase + 256, vm[0]);
#pragma unroll
            for (int c8 = 0; c8 < EPI_COLS / 8; ++c8) {
                const int cur = c8 & 1;
                tmem_ld_wait();
                if (c8 + 1 < EPI_COLS / 8) {                                   // next 8 columns in flight while these are scaled
                    tmem_ld8(tbase + 8 * (c8 + 1), vh[cur ^ 1]);
                    tmem_ld8(tbase + 128 + 8 * (c8 + 1), vl[cur ^ 1]);
                    tmem_ld8(tbase + 256 + 8 * (c8 + 1), vm[cur ^ 1]);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int sumi = ((int)vh[cur][i] << 3) + (int)vl[cur][i];
                    float t = __fmul_rn(dw.x, (float)sumi);
                    t = __fmaf_rn(-dw.y, (float)(int)vm[cur][i], t);
                    acc[8 * c8 + i] = __fmaf_rn(dxk[8 * c8 + i], t, acc[8 * c8 + i]);
                }
            }
            tc_fence_before();
            mbar_arrive(tempty_bar(h));
        }
        if (row < p.M) {
#pragma unroll
            for (int c = 0; c < EPI_COLS; ++c) {
                const int64_t n = n0 + EPI_COLS * grp + c;
                if (n < p.N) p.dst[n * p.dst_stride + row] = acc[c];
            }
        }
    } else if (warp < EPI_WARPS + PROD_WARPS) {
        // ===================== producers: Q4_K bytes -> three int8 A tiles; planar int8 activations -> B tile =====================
        const int t = threadIdx.x - EPI_WARPS * 32;
        const uint32_t smem0 = smem_addr(smem);
        const int64_t ntile = blockIdx.x;
#pragma unroll
        for (int i = 0; i < PREFETCH; ++i) {             // fresh slots: nothing to wait for
            if (i < nsteps) producer_issue(p, smem0, t, m0, ntile, i);
            cp_async_commit();
        }
        for (int step = 0; step < nsteps; ++step) {
            cp_async_wait<PREFETCH - 1>();               // this thread's copies for `step` have landed ...
            if ((step & 1) == 0) asm volatile("bar.sync 3, 128;" ::: "memory");   // ... and everybody else's: the raw rows are shared
            const int a = step % A_STAGES;
            mbar_wait(emptya_bar(a), (uint32_t)(((step / A_STAGES) & 1) ^ 1));
            producer_unpack(smem + SMEM_A + a * A_STAGE_BYTES, smem + SMEM_RAW + ((step >> 1) % RAW_SLOTS) * RAW_SLOT_BYTES + t * RAW_ROW_BYTES,
                            t, step & 1);
            fence_async_smem();                          // generic-proxy writes (st.shared, cp.async) -> tensor core's async proxy
            mbar_arrive(full_bar(step % B_SLOTS));
            const int ahead = step + PREFETCH;
            if (ahead < nsteps) {
                // the B slot of step `ahead` was last read by the MMAs of step ahead - B_SLOTS
                mbar_wait(emptyb_bar(ahead % B_SLOTS), (uint32_t)(((ahead / B_SLOTS) & 1) ^ 1));
                producer_issue(p, smem0, t, m0, ntile, ahead);
            }
            cp_async_commit();
        }
    } else if (lane == 0) {
        // ===================== MMA issuer (one thread) =====================
        const uint32_t idesc = umma_idesc_i8();
        const uint32_t stage0 = smem_addr(smem);
        for (int step = 0; step < nsteps; ++step) {
            const int kb = step >> 1, hh = step & 1, a = step % A_STAGES, bslot = step % B_SLOTS;
            mbar_wait(full_bar(bslot), (uint32_t)((step / B_SLOTS) & 1));
            tc_fence_after();
            const uint32_t sa = stage0 + SMEM_A + (uint32_t)a * A_STAGE_BYTES, sb = stage0 + SMEM_B + (uint32_t)bslot * TILE_BYTES;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (hh == 0) {                     // new superblock: the previous one must have left this half of TMEM
                    mbar_wait(tempty_bar(h), (uint32_t)((kb & 1) ^ 1));
                    tc_fence_after();
                }
#pragma unroll
                for (int o = 0; o < 3; ++o) {
#pragma unroll
                    for (int j = 0; j < BK / 32; ++j) {
                        const uint64_t a_desc = umma_desc(sa + o * TILE_BYTES + j * 2 * LBO_BYTES);
                        const uint64_t b_desc = umma_desc(sb + h * (HALF_N / 8) * SBO_BYTES + j * 2 * LBO_BYTES);
                        umma_i8(tmem + o * 128 + h * HALF_N, a_desc, b_desc, idesc, (uint32_t)((hh | j) != 0));
                    }
                }
                if (hh == 1) umma_commit(tfull_bar(h));     // superblock complete in this half -> its epilogue group
            }
            umma_commit(emptya_bar(a));                     // A stage and B slot consumed -> producers
            umma_commit(emptyb_bar(bslot));
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == EPI_WARPS + PROD_WARPS) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
    }
}

// F32 rows -> int8 activations in the GEMM's own tiled order + float d[nb][Nld]; the arithmetic of quantize_row_q8_K
// (shared device function, bit-exact), one warp per 256-block.  Element k of column n, superblock b lands in
//   tile (n / 128, b), half k / 128, then the canonical K-major core-matrix order of the shared-memory B tile:
//   ((n % 128) / 8) * 1024 + ((k % 128) / 16) * 128 + (n % 8) * 16 + k % 16
__global__ void __launch_bounds__(256)
quantize_q8_planar_kernel(const float *__restrict__ x, int8_t *__restrict__ qs, float *__restrict__ d, int64_t nb_per_row,
                          int64_t nblocks_total, int64_t x_row_stride_bytes, int64_t Nld) {
    const int lane = threadIdx.x & 31;
    const int64_t blk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (blk >= nblocks_total) return;
    const int64_t row = blk / nb_per_row, b = blk - row * nb_per_row;
    const float *src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + row * x_row_stride_bytes) + b * QK_K;
    const float4 a = __ldg(reinterpret_cast<const float4 *>(src) + lane);
    const float4 c = __ldg(reinterpret_cast<const float4 *>(src) + 32 + lane);
    const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
    uint32_t w_lo, w_hi;
    int s_lo, s_hi;
    const float dq = quantize_block_q8_K(v, lane, w_lo, w_hi, s_lo, s_hi);
    const int64_t ntile = row >> 7, nl = row & 127;
    int8_t *tile = qs + ((ntile * nb_per_row + b) << 15) + (nl >> 3) * 1024 + (nl & 7) * 16;
    // lane l holds elements 4l..4l+3 (half 0) and 128+4l..128+4l+3 (half 1): chunk l / 4, bytes 4 * (l % 4)
    const int off = (lane >> 2) * 128 + (lane & 3) * 4;
    *reinterpret_cast<uint32_t *>(tile + off) = w_lo;
    *reinterpret_cast<uint32_t *>(tile + 16384 + off) = w_hi;
    if (lane == 0) d[b * Nld + row] = dq;
}

}  // namespace

// Q8_K-block activations (q4k_mul_mat_q8 with N > 8): no planar copy is available without a workspace; the caller
// runs the decode kernel eight columns at a time (same GPU arithmetic)
cudaError_t launch_gemm_q4k_q8k(const void *, int64_t, int64_t, int64_t, const void *, int64_t, int64_t,
                                float *, int64_t, cudaStream_t) {
    return cudaErrorNotSupported;
}

// F32 activations: quantize into `workspace` (roundup(N,128)*K + nb*N*4 bytes <= q4k_mul_mat_workspace_bytes), then the
// tensor-core GEMM.  Two launches.
cudaError_t launch_gemm_q4k_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                const float *x, int64_t N, int64_t x_row_stride_bytes,
                                float *dst, int64_t dst_col_stride_bytes, void *workspace, cudaStream_t stream) {
    if (M == 0 || N == 0) return cudaSuccess;
    const int64_t nb = K / QK_K;
    const int64_t Npad = (N + BN - 1) / BN * BN;          // whole column tiles; columns >= N hold stale bytes, never stored
    int8_t *qs = reinterpret_cast<int8_t *>(workspace);
    float *d = reinterpret_cast<float *>(qs + Npad * K);
    const int64_t total = nb * N;
    quantize_q8_planar_kernel<<<(unsigned)((total + 7) / 8), 256, 0, stream>>>(x, qs, d, nb, total, x_row_stride_bytes, N);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;

    GemmParams p;
    p.w = (const uint8_t *)w; p.M = M; p.w_stride = w_row_stride_bytes; p.nb = (int)nb;
    p.xq = qs; p.xd = d; p.N = N; p.Nld = N;
    p.dst = dst; p.dst_stride = dst_col_stride_bytes / 4;
    constexpr size_t smem = SMEM_TOTAL;
    static bool attr_set = false;
    if (!attr_set) {
        e = cudaFuncSetAttribute(gemm_q4k_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const dim3 grid((unsigned)((N + BN - 1) / BN), (unsigned)((M + BM - 1) / BM));     // column tiles of one row tile run together (L2)
    gemm_q4k_tc_kernel<<<grid, GEMM_THREADS, smem, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}

}  // namespace q4k
