// gemm_q4k_tc.cu — prefill GEMM (batch > 8) on the 5th-generation tensor cores: tcgen05.mma, TMEM accumulators,
// weights fed to the tensor core from TENSOR MEMORY (the A operand never touches shared memory).
// dst[n][m] = sum_blk  d_x[n,blk] * ( d_w[m,blk] * sumi[m,n,blk]  -  dmin_w[m,blk] * summ[m,n,blk] )
// which is ggml_compute_forward_mul_mat for Q4_K x F32 at ne11 = N (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401) with
// the integer stage of ggml_vec_dot_q4_K_q8_K (L/ggml/src/ggml-cpu/quants.c:575-611) kept EXACT:
//   sumi = sum_j sc[j] * p[j]        p[j] = sum_{l<32} q4[32j+l] * q8[32j+l]
//   summ = sum_j m[j]  * bs[j]       bs[j] = sum of the 32 q8 of sub-block j
//
// sumi through an int8 tensor core, exactly.  sc <= 63 and q4 <= 15, so sc*q4 needs 10 bits.  Split the SCALE, not
// the product:  sc = 8*sh + sl  (sh, sl <= 7).  Then sh*q4 and sl*q4 are <= 105 and fit a signed byte, a packed
// multiply of four nibble-bytes by sh never carries between bytes, and  sumi = 8 * (A_hi . q8) + (A_lo . q8)  with
// A_hi = sh*q4, A_lo = sl*q4 accumulated over the whole 256-wide superblock in two int32 TMEM accumulators
// (kind::i8).
//
// summ through ONE small MMA per superblock, exactly.  bs[j] is in [-4064, 4064]; write bs = 64*bh + bl with
// bl = bs & 63, bh = bs >> 6 in [-64, 63].  Then  summ = sum_j (64 m[j]) * bh[j] + m[j] * bl[j]  is a K = 16 dot
// product whose factors (<= 4032, 6 significant bits) are exact in fp16, whose products are < 2^18 and whose sum is
// < 2^22: a kind::f16 MMA with an fp32 accumulator computes it without rounding.  The quantizer emits bh / bl.
//
// What bounds this kernel is shared-memory bandwidth and issue slots, not the tensor pipe (profiles/r1_gemm_trace_v4.txt:
// with A in shared memory the MMAs alone read 144 KB per K = 128 step, 1150 cycles at 128 B/clk).  So:
//   * A_hi / A_lo are written by the producer threads straight from registers into TMEM (tcgen05.st, thread <->
//     TMEM lane <-> weight row) and read from there by the MMA ([a_tmem] operand form);
//   * shared memory carries only the int8 activation tile (B), the raw Q4_K bytes and the tiny mins operands.
//
// CTA = 128 weight rows x 128 activation columns, 25 warps, warp-specialised:
//   warps 0-15   epilogue: thread <-> TMEM lane (= weight row) of quadrant warp % 4; column group g = warp / 4 owns
//                columns 16g..16g+15 of BOTH accumulator halves, so one half drains with all 16 warps while the
//                tensor core works on the other.
//   warps 16-23  producers: coalesced 16-byte cp.async of whole block_q4_K rows (raw ring), of the pre-tiled int8
//                activation tile (straight into its B slot) and of the mins tile; then thread <-> weight row
//                (quadrant = warp % 4) and K half (warp / 4): unpack scales, multiply nibbles, tcgen05.st.
//   warp 24      allocates TMEM; one lane issues the MMAs (M = 128, N = 64, K = 32) and commits to mbarriers.
// TMEM columns: [0,128) A_hi.q8, [128,256) A_lo.q8, [256,384) mins (fp32), [384,512) two A stages x (hi, lo) x 32.
// The two 64-column halves of the accumulators alternate between the tensor core and the epilogue.
//
// Activations (quantize_q8_planar_kernel below): the same int8 / float d that quantize_row_q8_K produces (shared
// device function, bit-exact), int8 in tile order + float d[nb][N] + the fp16 mins tile (bh | bl).
#include <algorithm>
#include <cstdlib>

#include <cuda.h>
#include <cuda_fp16.h>

#include "gemm_tc_ptx.cuh"
#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

constexpr int BN = 128, BK = 128;                    // CTA tile (BM = 128 rows); BK = K bytes per pipeline step (half a superblock)
constexpr int B_SLOTS = 6;                            // activation tiles, written in place by cp.async
constexpr int RAW_SLOTS = 4;                          // raw block_q4_K bytes of 128 rows x one superblock
constexpr int AM_SLOTS = 2, BMIN_SLOTS = 4;           // mins operands (per superblock): weights side, activation side
constexpr int TILE_BYTES = BN * BK;                   // 16 KB: the int8 activation tile of a step
constexpr int MIN_TILE_BYTES = 128 * 16 * 2;          // 4 KB: 128 rows (or columns) x 16 fp16
constexpr int RAW_ROW_BYTES = BLK_Q4K, RAW_SLOT_BYTES = BM * RAW_ROW_BYTES;
static_assert(B_SLOTS % 2 == 0 && AM_SLOTS == 2 && (BMIN_SLOTS & (BMIN_SLOTS - 1)) == 0, "the MMA issuer's running slot indices");
constexpr int EPI_WARPS = 16, PROD_WARPS = 8;
constexpr int PROD_THREADS = PROD_WARPS * 32;
constexpr int EPI_COLS = 16;                          // columns per epilogue thread per half
constexpr int MMA_WARP = EPI_WARPS + PROD_WARPS, LOAD_WARP = MMA_WARP + 1;
constexpr int GEMM_THREADS = (LOAD_WARP + 1) * 32;   // 832
constexpr int SMEM_B = 0, SMEM_RAW = SMEM_B + B_SLOTS * TILE_BYTES, SMEM_AM = SMEM_RAW + RAW_SLOTS * RAW_SLOT_BYTES;
constexpr int SMEM_BMIN = SMEM_AM + AM_SLOTS * MIN_TILE_BYTES, SMEM_DX = SMEM_BMIN + BMIN_SLOTS * MIN_TILE_BYTES;
constexpr int DX_SLOTS = 8, DX_BYTES = BN * 4;        // d_x of the tile's 128 columns for one superblock, copied by the loader with the B tile
constexpr int SMEM_BARS = SMEM_DX + DX_SLOTS * DX_BYTES;
constexpr int SMEM_TOTAL = SMEM_BARS + 256;
constexpr int TMEM_COLS = 512;
constexpr uint32_t TM_HI = 0, TM_LO = 128, TM_MIN = 256, TM_A = 384;   // TMEM column of each region
constexpr uint32_t TM_A_STAGE = 64, TM_A_TILE = 32;  // columns per A stage (hi + lo) and per tile (128 bytes of K)
constexpr uint32_t SBO_BYTES = 8 * BK;                // int8 B tile: next 8-column group of core matrices
constexpr uint32_t MIN_SBO_BYTES = 256;               // mins tiles: 8 rows x (2 chunks x 16 bytes)
constexpr size_t FLAG_BYTES = 1024;                   // one flag word per CTA behind the quantized activations
constexpr size_t MAX_CTAS = 160, PARTIAL_BYTES = (size_t)BM * 128 * 4;   // persistent grid <= SM count; a parked 128 x 128 fp32 tile per CTA

#ifdef Q4K_GEMM_TRACE
__device__ long long g_gemm_trace[16][256];
#define TR(slot, idx) do { if (blockIdx.x == 0 && blockIdx.y == 0 && (idx) < 256) g_gemm_trace[slot][idx] = clock64(); } while (0)
#else
#define TR(slot, idx) do { } while (0)
#endif

struct GemmParams {
    // up to four weight matrices with the same K that read the SAME activations (attn_q / attn_k / attn_v, ffn_gate / ffn_up) in one
    // launch: row tiles are numbered through all of them, tend[i] = row tiles of matrices 0..i
    CUtensorMap wmap[4];    // the weights as 2-D byte tensors [M][row stride], box = 128 rows x one block_q4_K (144 bytes)
    const uint8_t *w[4];
    int64_t M[4];
    float *dst[4];
    int64_t dst_stride[4];  // floats between dst columns
    int tend[4], nmat;
    int64_t w_stride;
    int nb;
    const int8_t *xq;       // int8 activations, pre-tiled: [column tile][superblock][half][16 KB canonical B tile]
    const float *xd;        // [nb][Nld]
    const uint8_t *bmin;    // fp16 mins tiles: [column tile][superblock][4 KB canonical tile: 128 columns x (bh[8] | bl[8])]
    int64_t N, Nld;
    int ntn;                // column tiles
    int halves;             // 1: the work units are 64-column HALF tiles (few tiles, every tile has two real halves: twice the CTAs, each half as wide)
    int64_t units;          // (tile, superblock) work units of the launch = tiles x nb, cut into gridDim.x equal ranges
    int *flags;             // one zero-initialised word per CTA: "this CTA's parked partial sums are in the workspace"
    float *partial;         // [gridDim.x][128 columns][128 rows] fp32: the sums of a range's first tile when that tile started earlier
    int32_t *dump_sumi, *dump_summ;   // DUMP instantiation (parity): [(n*M + m)*nb + b] = the integers the epilogue read from TMEM
};

// Global -> shared copies of one pipeline step, issued by one lane of the loader warp:
//   B tile of `step`        : one 16 KB bulk copy (the activations are stored pre-tiled in exactly the shared-memory order)
//   mins tile of `sb`       : one 4 KB bulk copy                                          } with the B tile of step 2*sb
//   raw weights of `sb`     : one TMA box, 128 rows x 144 bytes (rows past M read as zero) }
// Slots are indexed by the CTA's running step / superblock count (they run on through the CTA's work items).
// The raw weights do not depend on the quantizer kernel in front of this one (programmatic dependent launch), the B and
// mins tiles do: the loader issues the first raw boxes before griddepcontrol.wait.
__device__ __forceinline__ void loader_issue_raw(const CUtensorMap *wmap, uint32_t smem0, uint32_t rawfull, int m0, int kb, int kb0) {
    mbar_expect_tx(rawfull, RAW_SLOT_BYTES);
    tma_g2s_2d(smem0 + SMEM_RAW + (kb % RAW_SLOTS) * RAW_SLOT_BYTES, wmap, (kb0 + kb) * BLK_Q4K, m0, rawfull);
}
// the thread's 64 weights of this step (one 32-byte group of its raw row: low nibbles = sub-block 2j, high nibbles
// = sub-block 2j + 1) -> 2 x 16 TMEM columns: A_hi = sh * q4, A_lo = sl * q4, K order = element order
__device__ __forceinline__ void producer_unpack(uint32_t (&hi)[16], uint32_t (&lo)[16], const uint8_t *raw, const uint4 sm, int j) {
    const uint4 qa = *reinterpret_cast<const uint4 *>(raw + 16 + 32 * j), qb = *reinterpret_cast<const uint4 *>(raw + 32 + 32 * j);
    const uint32_t qw[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
    const uint32_t sc2 = byte_pair(sm.x, sm.y, j);
#pragma unroll
    for (int nib = 0; nib < 2; ++nib) {
        const uint32_t sc = (sc2 >> (8 * nib)) & 0xffu, sh = sc >> 3, sl = sc & 7u;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t v = (nib ? (qw[i] >> 4) : qw[i]) & 0x0f0f0f0fu;
            hi[8 * nib + i] = v * sh;
            lo[8 * nib + i] = v * sl;
        }
    }
}

// the row's mins operand of a superblock: 64 * m[0..7] | m[0..7] as fp16, canonical K-major tile (SBO 256, LBO 128)
__device__ __forceinline__ void producer_mins(uint8_t *am_tile, int row, const uint4 sm) {
    const __half2 bias = __half2half2(__ushort_as_half((unsigned short)0x6400)), k64 = __half2half2(__ushort_as_half((unsigned short)0x5400));
    uint32_t m1[4], m64[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t src = i < 2 ? sm.z : sm.w;
        const uint32_t pair = __byte_perm(src, 0u, (i & 1) ? 0x4342u : 0x4140u) | 0x64006400u;        // {1024 + m[2i], 1024 + m[2i+1]}
        const __half2 h = __hsub2(*reinterpret_cast<const __half2 *>(&pair), bias);                   // exact
        const __half2 h64 = __hmul2(h, k64);                                                          // exact (<= 4032, 6 bits)
        m1[i] = *reinterpret_cast<const uint32_t *>(&h);
        m64[i] = *reinterpret_cast<const uint32_t *>(&h64);
    }
    uint8_t *dst = am_tile + (row >> 3) * MIN_SBO_BYTES + (row & 7) * 16;
    *reinterpret_cast<uint4 *>(dst) = make_uint4(m64[0], m64[1], m64[2], m64[3]);
    *reinterpret_cast<uint4 *>(dst + LBO_BYTES) = make_uint4(m1[0], m1[1], m1[2], m1[3]);
}

// One work item of a CTA: superblocks [kb0, kb0 + nbk) of output tile (mat, m0, n0).  The launch is PERSISTENT and split
// "stream-K" fashion: all (tile, superblock) units of the launch — tiles in (row tile, column tile) order, column tiles of one
// row tile next to each other (L2) — form one sequence that is cut into gridDim.x contiguous, equal ranges, so every SM works
// the same number of superblocks whatever the tile count (448 tiles on 148 SMs used to be four waves for three waves' work,
// the 64-80 tiles of TinyLlama's matrices left half the SMs idle).  A range covers the tail of one tile, whole tiles, and the
// head of another.  Every role of the CTA walks the same item sequence; barrier slots and phases run on through the items
// (counters, not tile-local indices), so the next tile's weights and activations are in flight while this one's last
// accumulators drain.
struct WorkItem {
    int tile, kb0, nbk;                 // what the loops need: everything else is derived from `tile` where it is used (registers)
    int mat, h0, h1, m0, n0, ntile;     // [h0, h1): the 64-column halves of the tile this item computes
    __device__ __forceinline__ void decode(const GemmParams &p) {
        const int t = p.halves ? tile >> 1 : tile;
        const int rt = t / p.ntn;
        ntile = t - rt * p.ntn;
        mat = 0;
        while (mat + 1 < p.nmat && rt >= p.tend[mat]) ++mat;
        m0 = (rt - (mat ? p.tend[mat - 1] : 0)) * BM;
        n0 = ntile * BN;
        if (p.halves) { h0 = tile & 1; h1 = h0 + 1; }
        else { h0 = 0; h1 = (int64_t)n0 + HALF_N < p.N ? 2 : 1; }     // halves that hold real columns
    }
};

struct WorkRange {
    int u, u1;                          // units fit 32 bits (the launcher checks)
    __device__ __forceinline__ WorkRange(const GemmParams &p, int64_t cta, int64_t nctas) : u((int)(p.units * cta / nctas)), u1((int)(p.units * (cta + 1) / nctas)) {}
    __device__ __forceinline__ bool next(const GemmParams &p, WorkItem &it) {
        if (u >= u1) return false;
        it.tile = u / p.nb;
        it.kb0 = u - it.tile * p.nb;
        it.nbk = min(p.nb - it.kb0, u1 - u);
        u += it.nbk;
        it.decode(p);
        return true;
    }
};

// The epilogue's superblock loop for one work item.  HS: 0 / 1 = only that 64-column half, 2 = both.  Everything in here is on the
// kernel's critical path (the tensor core waits for its accumulator halves): run time follows this loop's instruction count.
template <bool DUMP, int HS>
__device__ __forceinline__ void drain_superblocks(const GemmParams &p, const WorkItem &it, uint8_t *smem, uint32_t bar0, uint32_t tbase,
                                                  int qd, int g, int lane, uint32_t &hpar, uint32_t &gsb, uint64_t (&acc)[2][EPI_COLS / 2]) {
    for (int kb = 0; kb < it.nbk; ++kb) {
        // {d, dmin} of this row's block: the first four bytes of its raw row, in the loader's slot since before the producers
        // unpacked it (the wait below is long complete: it acquires the TMA's bytes) and until every epilogue thread has
        // handed this superblock's accumulators back (the loader refills the slot for superblock + 4 behind MMAs that need them)
        mbar_wait(bar0 + 8u * (2 * B_SLOTS + 4 + (gsb % RAW_SLOTS)), (gsb / RAW_SLOTS) & 1u);
        const uint32_t dd = *reinterpret_cast<const uint32_t *>(smem + SMEM_RAW + (gsb % RAW_SLOTS) * RAW_SLOT_BYTES + (32 * qd + lane) * RAW_ROW_BYTES);
        const float2 dw = __half22float2(*reinterpret_cast<const __half2 *>(&dd));
        const uint64_t d2 = pack2(dw.x, dw.x), ndmin2 = pack2(-dw.y, -dw.y);
#pragma unroll
        for (int h = (HS == 1 ? 1 : 0); h < (HS == 0 ? 1 : 2); ++h) {
            mbar_wait_long<Q4K_GEMM_WAIT_TFULL>(bar0 + 8u * (2 * B_SLOTS + h), (hpar >> h) & 1u);
            hpar ^= 1u << h;
            tc_fence_after();
            // (d_x arrived on the step's `full` barrier, which the MMAs behind this tfull had waited for)
            const uint64_t *dxk = reinterpret_cast<const uint64_t *>(smem + SMEM_DX + (gsb % DX_SLOTS) * DX_BYTES + (HALF_N * h + EPI_COLS * g) * 4);
#pragma unroll
            for (int c8 = 0; c8 < EPI_COLS / 8; ++c8) {
                uint32_t vh[8], vl[8], vm[8];
                const uint32_t ta = tbase + HALF_N * h + 8 * c8;
                tmem_ld8(ta + TM_HI, vh); tmem_ld8(ta + TM_LO, vl); tmem_ld8(ta + TM_MIN, vm);
                tmem_ld_wait();
                if (c8 == EPI_COLS / 8 - 1) {          // everything of this half is in registers: hand the columns back
                    tc_fence_before();
                    mbar_arrive(bar0 + 8u * (2 * B_SLOTS + 2 + h));
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {     // t = d * sumi - dmin * summ (two roundings), acc = fma(d_x, t, acc): pairs of columns
                    const int s0 = ((int)vh[2 * i] << 3) + (int)vl[2 * i], s1 = ((int)vh[2 * i + 1] << 3) + (int)vl[2 * i + 1];
                    if constexpr (DUMP) {
                        const int64_t M_mat = p.M[it.mat], row = (int64_t)it.m0 + 32 * qd + lane;
                        const int64_t n = (int64_t)it.n0 + HALF_N * h + EPI_COLS * g + 8 * c8 + 2 * i;
                        if (row < M_mat && n < p.N) {
                            p.dump_sumi[(n * M_mat + row) * p.nb + it.kb0 + kb] = s0;
                            p.dump_summ[(n * M_mat + row) * p.nb + it.kb0 + kb] = (int)__uint_as_float(vm[2 * i]);
                        }
                        if (row < M_mat && n + 1 < p.N) {
                            p.dump_sumi[((n + 1) * M_mat + row) * p.nb + it.kb0 + kb] = s1;
                            p.dump_summ[((n + 1) * M_mat + row) * p.nb + it.kb0 + kb] = (int)__uint_as_float(vm[2 * i + 1]);
                        }
                    }
                    // (packed f32x2: measured faster than scalar fp32 despite the register-pair moves it needs)
                    uint64_t t = mul2(d2, pack2((float)s0, (float)s1));
                    t = fma2(ndmin2, pack2(__uint_as_float(vm[2 * i]), __uint_as_float(vm[2 * i + 1])), t);
                    acc[h][4 * c8 + i] = fma2(dxk[4 * c8 + i], t, acc[h][4 * c8 + i]);
                }
            }
        }
        ++gsb;
    }
}

// DUMP: parity instantiation — the same code; the epilogue also writes out every superblock's sumi (= 8 hi + lo) and summ
template <bool DUMP>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_q4k_tc_kernel(const __grid_constant__ GemmParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + SMEM_BARS);
    uint32_t *tmem_ptr_s = reinterpret_cast<uint32_t *>(bars + 24);
    const uint32_t bar0 = smem_addr(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };                  // [B_SLOTS]  producers -> MMA, one per pipeline step slot
    auto done_bar = [&](int s) { return bar0 + 8u * (B_SLOTS + s); };      // [B_SLOTS]  MMA -> producers: the step's MMAs have completed
    auto tfull_bar = [&](int h) { return bar0 + 8u * (2 * B_SLOTS + h); };
    auto tempty_bar = [&](int h) { return bar0 + 8u * (2 * B_SLOTS + 2 + h); };
    auto rawfull_bar = [&](int s) { return bar0 + 8u * (2 * B_SLOTS + 4 + s); };   // [RAW_SLOTS] loader -> producers

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (threadIdx.x == 0) {
        for (int s = 0; s < B_SLOTS; ++s) { mbar_init(full_bar(s), PROD_THREADS + 1); mbar_init(done_bar(s), 1); }
        for (int s = 0; s < RAW_SLOTS; ++s) mbar_init(rawfull_bar(s), 1);
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
        WorkRange range(p, blockIdx.x, gridDim.x);
        WorkItem it;
        // ===================== epilogue: TMEM -> registers, per-superblock fp32 scales =====================
        const int qd = warp & 3, g = warp >> 2;                               // lane quadrant, column group (16 wide, both halves)
        // The scales of a superblock come out of shared memory: d_x of the tile's 128 columns is copied by the loader next to the B tile
        // (a ring of 8: refilled only behind MMAs that need this superblock's accumulators back twice over), {d, dmin} of a row are
        // the first four bytes of its raw row in the loader's slot.  In this loop every instruction counts: run time follows the
        // epilogue's instruction count almost one to one (profiles/r2_gemm_streamk.txt).  Tried and slower: plain global loads one
        // superblock ahead (sunk to their use by the compiler), cp.async per lane (31 extra shared-memory wavefronts for the rows).
        // (warp-uniform: the shuffle lets the compiler keep it in a uniform register instead of spilling it around the loop)
        const uint32_t tbase = __shfl_sync(0xffffffffu, tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(EPI_COLS * g), 0);
        uint32_t hpar = 0u;                                                   // bit h: phase of accumulator half h's barriers
        uint32_t gsb = 0u;                                                    // superblocks of this CTA so far (the loader's slots)
        while (range.next(p, it)) {
            uint64_t acc[2][EPI_COLS / 2];                                        // fp32 pairs
#pragma unroll
            for (int i = 0; i < EPI_COLS / 2; ++i) acc[0][i] = acc[1][i] = 0ull;
            // the superblock loop, specialised for "both halves" (the common case: no test, no merge of the accumulators behind a branch —
            // that alone cost 16 register moves per superblock) and for one half (ragged last column tile, half-tile units)
            if (it.h0 == 0 && it.h1 == 2) drain_superblocks<DUMP, 2>(p, it, smem, bar0, tbase, qd, g, lane, hpar, gsb, acc);
            else if (it.h0 == 0) drain_superblocks<DUMP, 0>(p, it, smem, bar0, tbase, qd, g, lane, hpar, gsb, acc);
            else drain_superblocks<DUMP, 1>(p, it, smem, bar0, tbase, qd, g, lane, hpar, gsb, acc);
            // ---- the item's sums leave the registers.  A tile that lies inside one range is stored.  A split tile: the CTAs that
            // hold its LATER superblocks meet it FIRST in their ranges and park their sums in the workspace (one slot per CTA, then
            // a flag); the CTA that holds its first superblocks meets it LAST, adds the parked parts in K order and stores — it
            // never waits for work that has not been done long ago, and nobody waits for it.  One fixed association per launch
            // shape and SM count; two runs give the same bits.
            {   // the tile's coordinates again, from the tile number alone: nothing of them stays live through the superblock loop
                int tile = it.tile;
                asm volatile("" : "+r"(tile));
                it.tile = tile;
                it.decode(p);
            }
            const int64_t M_mat = p.M[it.mat];
            const int64_t row = (int64_t)it.m0 + 32 * qd + lane;
            const bool first = it.kb0 == 0, last = it.kb0 + it.nbk == p.nb;
            if (!first) {
                float *part = p.partial + (size_t)blockIdx.x * (BM * BN);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (h < it.h0 || h >= it.h1) continue;
#pragma unroll
                    for (int c = 0; c < EPI_COLS / 2; ++c) {
                        float a0, a1;
                        unpack2(acc[h][c], a0, a1);
                        const int col = HALF_N * h + EPI_COLS * g + 2 * c;
                        __stcg(part + col * BM + 32 * qd + lane, a0);
                        __stcg(part + (col + 1) * BM + 32 * qd + lane, a1);
                    }
                }
                __threadfence();
                asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");      // all 16 epilogue warps have stored
                if (threadIdx.x == 0) asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p.flags + blockIdx.x), "r"(1) : "memory");
            } else {
                if (!last) {
                    // the parts that follow: CTA c + 1 starts with superblock kb0 + nbk of this tile, and so on until K is covered
                    int covered = it.nbk;
                    for (int64_t c = (int64_t)blockIdx.x + 1; covered < p.nb; ++c) {
                        const int64_t r0 = p.units * c / gridDim.x, r1 = p.units * (c + 1) / gridDim.x;
                        if (lane == 0) {
                            int seen;
                            do { asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(p.flags + c) : "memory"); } while (seen != 1);
                        }
                        __syncwarp();
                        const float *part = p.partial + (size_t)c * (BM * BN);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            if (h < it.h0 || h >= it.h1) continue;
                            float prev[EPI_COLS];
#pragma unroll
                            for (int cc = 0; cc < EPI_COLS; ++cc) prev[cc] = __ldcg(part + (HALF_N * h + EPI_COLS * g + cc) * BM + 32 * qd + lane);
#pragma unroll
                            for (int cc = 0; cc < EPI_COLS / 2; ++cc) {
                                float a0, a1;
                                unpack2(acc[h][cc], a0, a1);
                                acc[h][cc] = pack2(__fadd_rn(a0, prev[2 * cc]), __fadd_rn(a1, prev[2 * cc + 1]));
                            }
                        }
                        covered += (int)min((int64_t)(p.nb - covered), r1 - r0);
                    }
                }
                if (row < M_mat) {
                    float *dst_mat = p.dst[it.mat];
                    const int64_t dst_stride = p.dst_stride[it.mat];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if (h < it.h0 || h >= it.h1) continue;
#pragma unroll
                        for (int c = 0; c < EPI_COLS / 2; ++c) {
                            const int64_t n = (int64_t)it.n0 + HALF_N * h + EPI_COLS * g + 2 * c;
                            float a0, a1;
                            unpack2(acc[h][c], a0, a1);
                            if (n < p.N) dst_mat[n * dst_stride + row] = a0;
                            if (n + 1 < p.N) dst_mat[(n + 1) * dst_stride + row] = a1;
                        }
                    }
                }
            }
        }
    } else if (warp < EPI_WARPS + PROD_WARPS) {
        // ===================== producers: Q4_K bytes -> A_hi / A_lo in TMEM; activations -> B slot =====================
        const int t = threadIdx.x - EPI_WARPS * 32, pw = t >> 5;
        const int qd = pw & 3, khalf = pw >> 2, r = 32 * qd + lane;           // TMEM lane = weight row of the tile; which 64 of the 128 K
        const uint32_t ta_base = __shfl_sync(0xffffffffu, tmem + ((uint32_t)(32 * qd) << 16) + TM_A + 16u * khalf, 0);
        uint4 sm = make_uint4(0, 0, 0, 0);
        int gstep = 0;                                                        // pipeline steps of this CTA so far
        WorkRange range(p, blockIdx.x, gridDim.x);
        WorkItem it;
        while (range.next(p, it)) {
            for (int ls = 0; ls < 2 * it.nbk; ++ls, ++gstep) {
                const int gkb = gstep >> 1, hh = gstep & 1;
                if (hh == 0) mbar_wait_long<Q4K_GEMM_WAIT>(rawfull_bar(gkb % RAW_SLOTS), (uint32_t)((gkb / RAW_SLOTS) & 1));
                const uint8_t *raw = smem + SMEM_RAW + (gkb % RAW_SLOTS) * RAW_SLOT_BYTES + r * RAW_ROW_BYTES;
                if (hh == 0) sm = unpack_scales_mins(*reinterpret_cast<const uint4 *>(raw));
                uint32_t hi[16], lo[16];
                producer_unpack(hi, lo, raw, sm, 2 * hh + khalf);
                // A stage gstep % 2 was last read by the MMAs of step gstep - 2
                if (gstep >= 2) mbar_wait_long<Q4K_GEMM_WAIT>(done_bar((gstep - 2) % B_SLOTS), (uint32_t)(((gstep - 2) / B_SLOTS) & 1));
                tc_fence_after();
                const uint32_t ta = ta_base + (uint32_t)(gstep & 1) * TM_A_STAGE;
                tmem_st16(ta, hi);
                tmem_st16(ta + TM_A_TILE, lo);
                if (hh == 0 && khalf == 0) {
                    producer_mins(smem + SMEM_AM + (gkb % AM_SLOTS) * MIN_TILE_BYTES, r, sm);
                    fence_async_smem();                      // generic-proxy st.shared -> the tensor core's async proxy
                }
                tmem_st_wait();
                tc_fence_before();
                mbar_arrive(full_bar(gstep % B_SLOTS));
            }
        }
    } else if (warp == MMA_WARP) {
        // ===================== MMA issuer: the whole warp runs the loop (uniform control flow keeps the descriptors in
        // uniform registers), one elected lane issues =====================
        const uint32_t idesc = umma_idesc(false), idesc_min = umma_idesc(true);
        const uint32_t smem0 = smem_addr(smem);
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
        const uint64_t desc_hi = umma_desc(0, SBO_BYTES), desc_min_hi = umma_desc(0, MIN_SBO_BYTES);
        uint32_t hpar = 0u;
        // running slot indices instead of step % 6 etc. (this warp's instruction count sits between the MMAs): B slots are taken
        // two per superblock (6 slots: a superblock's pair never wraps), the A stage is the K half, mins slots follow the superblock
        int bslot = 0, amslot = 0, bmslot = 0;
        uint32_t bpar = 0u;
        WorkRange range(p, blockIdx.x, gridDim.x);
        WorkItem it;
        // Per superblock: all 16 MMAs of column half 0 (both K steps), then all 16 of half 1.  A half's TMEM columns are
        // drained by the epilogue while the tensor core works on the other half (a window of half a superblock).
        while (range.next(p, it)) {
            const int h0 = it.h0, h1 = it.h1;
            for (int kb = 0; kb < it.nbk; ++kb) {
                const uint64_t a_min = desc_min_hi + (uint64_t)((smem0 + SMEM_AM + amslot * MIN_TILE_BYTES) >> 4);
                const uint64_t b_min = desc_min_hi + (uint64_t)((smem0 + SMEM_BMIN + bmslot * MIN_TILE_BYTES) >> 4);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (h < h0 || h >= h1) continue;
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        if (h == h0) mbar_wait_long<Q4K_GEMM_WAIT>(full_bar(bslot + hh), bpar);   // the step's operands
                        if (hh == 0) {                 // new superblock: the previous one must have left this half of TMEM
                            mbar_wait_long<Q4K_GEMM_WAIT_TEMPTY>(tempty_bar(h), ((hpar >> h) & 1u) ^ 1u);
                            hpar ^= 1u << h;
                        }
                        tc_fence_after();
                        const uint32_t ta = tm + TM_A + (uint32_t)hh * TM_A_STAGE;
                        const uint64_t bd0 = desc_hi + (uint64_t)((smem0 + SMEM_B + (uint32_t)(bslot + hh) * TILE_BYTES) >> 4);
                        if (elect_one()) {
#pragma unroll
                            for (int o = 0; o < 2; ++o) {
#pragma unroll
                                for (int j = 0; j < BK / 32; ++j) {
                                    const uint64_t b_desc = bd0 + (uint64_t)((h * (HALF_N / 8) * SBO_BYTES + j * 2 * LBO_BYTES) >> 4);
                                    umma_i8_ts(tm + (o ? TM_LO : TM_HI) + h * HALF_N, ta + o * TM_A_TILE + 8 * j, b_desc, idesc, (uint32_t)((hh | j) != 0));
                                }
                            }
                            if (hh == 1) {             // the superblock's mins (K = 16, fp16), then this half is complete -> epilogue
                                umma_f16_ss(tm + TM_MIN + h * HALF_N, a_min, b_min + (uint64_t)((h * (HALF_N / 8) * MIN_SBO_BYTES) >> 4), idesc_min, 0u);
                                umma_commit(tfull_bar(h));
                            }
                            if (h == h1 - 1) umma_commit(done_bar(bslot + hh));   // A stage and B slot consumed by every half -> producers
                        }
                        __syncwarp();
                    }
                }
                bslot += 2;
                if (bslot == B_SLOTS) { bslot = 0; bpar ^= 1u; }
                amslot ^= 1;
                bmslot = (bmslot + 1) & (BMIN_SLOTS - 1);
            }
        }
    } else if (lane == 0) {
        // ===================== loader (one thread): TMA copies, up to B_SLOTS steps ahead of the tensor core — through the item
        // boundaries: the next tile's operands are on their way while this one's last steps run =====================
        const uint32_t smem0 = smem_addr(smem);
        int gstep = 0, pre = 0;
        bool waited = false;
        WorkRange range(p, blockIdx.x, gridDim.x);
        WorkItem it;
        while (range.next(p, it)) {
            const CUtensorMap *wmap = &p.wmap[it.mat];
            if (!waited) {                                          // raw boxes that can go out before the quantizer has finished
                pre = min(it.nbk, RAW_SLOTS);
                for (int kb = 0; kb < pre; ++kb) loader_issue_raw(wmap, smem0, rawfull_bar(kb % RAW_SLOTS), (int)it.m0, kb, it.kb0);
                asm volatile("griddepcontrol.wait;" ::: "memory");  // the activation tiles come from the kernel in front of this one
                waited = true;
            }
            for (int ls = 0; ls < 2 * it.nbk; ++ls, ++gstep) {
                // slot reuse: the MMAs of step - B_SLOTS have completed, so its B slot is free; every producer had arrived for that
                // step, so the raw slot of superblock step / 2 - RAW_SLOTS has been read; the mins slots likewise
                if (gstep >= B_SLOTS) mbar_wait_long<Q4K_GEMM_WAIT>(done_bar(gstep % B_SLOTS), (uint32_t)(((gstep - B_SLOTS) / B_SLOTS) & 1));
                const int gkb = gstep >> 1;
                if ((gstep & 1) == 0 && gkb >= pre) {
                    mbar_expect_tx(rawfull_bar(gkb % RAW_SLOTS), RAW_SLOT_BYTES);
                    tma_g2s_2d(smem0 + SMEM_RAW + (gkb % RAW_SLOTS) * RAW_SLOT_BYTES, wmap, (it.kb0 + (ls >> 1)) * BLK_Q4K, (int)it.m0, rawfull_bar(gkb % RAW_SLOTS));
                }
                const int hh = gstep & 1, akb = it.kb0 + (ls >> 1);
                const uint32_t full = full_bar(gstep % B_SLOTS);
                // half-tile units copy only their 64 columns: the first or second 8 KB of the B tile, 2 KB of the mins tile
                const uint32_t bsz = p.halves ? TILE_BYTES / 2 : TILE_BYTES, msz = p.halves ? MIN_TILE_BYTES / 2 : MIN_TILE_BYTES;
                const uint32_t boff = p.halves ? it.h0 * bsz : 0u, moff = p.halves ? it.h0 * msz : 0u;
                mbar_expect_tx(full, bsz + (hh == 0 ? msz + DX_BYTES : 0));
                if (hh == 0) bulk_g2s(smem0 + SMEM_DX + (gkb % DX_SLOTS) * DX_BYTES, p.xd + (int64_t)akb * p.Nld + it.n0, DX_BYTES, full);
                bulk_g2s(smem0 + SMEM_B + (gstep % B_SLOTS) * TILE_BYTES + boff, p.xq + ((((int64_t)it.ntile * p.nb + akb) * 2 + hh) << 14) + boff, bsz, full);
                if (hh == 0) bulk_g2s(smem0 + SMEM_BMIN + (gkb % BMIN_SLOTS) * MIN_TILE_BYTES + moff, p.bmin + (((int64_t)it.ntile * p.nb + akb) << 12) + moff, msz, full);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
    }
}

// F32 rows -> int8 activations in the GEMM's own tiled order + float d[nb][Nld] + the fp16 mins tile; the arithmetic
// of quantize_row_q8_K (shared device function, bit-exact), one warp per 256-block.  Element k of column n,
// superblock b lands in
//   tile (n / 128, b), half k / 128, then the canonical K-major core-matrix order of the shared-memory B tile:
//   ((n % 128) / 8) * 1024 + ((k % 128) / 16) * 128 + (n % 8) * 16 + k % 16
// and the sums bs[j] of its eight 32-element sub-blocks, split bs = 64 * bh + bl (bl = bs & 63), land as fp16 in
//   mins tile (n / 128, b):  ((n % 128) / 8) * 256 + {bh: 0, bl: 128} + (n % 8) * 16 + 2 * j
__global__ void __launch_bounds__(256)
quantize_q8_planar_kernel(const float *__restrict__ x, int8_t *__restrict__ qs, float *__restrict__ d, uint8_t *__restrict__ bmin,
                          int64_t nb_per_row, int64_t nblocks_total, int64_t x_row_stride_bytes, int64_t Nld, int *flags) {
    // programmatic dependent launch on both sides: this kernel overwrites the workspace the GEMM in front of it may still be
    // reading, so everything waits for that; the GEMM behind it may start its prologue (barriers, TMEM, raw weights) now
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (flags && blockIdx.x == 0) flags[threadIdx.x] = 0;        // the per-CTA "parked tile" flags of the GEMM that follows (FLAG_BYTES = 256 words)
    const int lane = threadIdx.x & 31;
    const int64_t blk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (blk >= nblocks_total) return;
    const int64_t row = blk / nb_per_row, b = blk - row * nb_per_row;
    const float *src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + row * x_row_stride_bytes) + b * QK_K;
    // plain (coherent) loads: under programmatic dependent launch the kernel that produces x may still have been running at entry
    const float4 a = ld_act16(src + 4 * lane);
    const float4 c = ld_act16(src + 128 + 4 * lane);
    const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
    uint32_t w_lo, w_hi;
    int s_lo, s_hi;
    const float dq = quantize_block_q8_K<8>(v, lane, w_lo, w_hi, s_lo, s_hi);     // 8-lane sums = the 32-element sub-block sums
    const int64_t ntile = row >> 7, nl = row & 127;
    int8_t *tile = qs + ((ntile * nb_per_row + b) << 15) + (nl >> 3) * 1024 + (nl & 7) * 16;
    // lane l holds elements 4l..4l+3 (half 0) and 128+4l..128+4l+3 (half 1): chunk l / 4, bytes 4 * (l % 4)
    const int off = (lane >> 2) * 128 + (lane & 3) * 4;
    *reinterpret_cast<uint32_t *>(tile + off) = w_lo;
    *reinterpret_cast<uint32_t *>(tile + 16384 + off) = w_hi;
    if (lane == 0) d[b * Nld + row] = dq;
    if ((lane & 7) == 0) {                                  // sub-blocks lane / 8 (half 0) and 4 + lane / 8 (half 1)
        __half *mt = reinterpret_cast<__half *>(bmin + ((ntile * nb_per_row + b) << 12) + (nl >> 3) * MIN_SBO_BYTES + (nl & 7) * 16);
        const int j = lane >> 3;
        mt[j] = __int2half_rn(s_lo >> 6);        mt[64 + j] = __int2half_rn(s_lo & 63);
        mt[4 + j] = __int2half_rn(s_hi >> 6);    mt[64 + 4 + j] = __int2half_rn(s_hi & 63);
    }
}

}  // namespace

// Q8_K-block activations (q4k_mul_mat_q8 with N > 8): no planar copy is available without a workspace; the caller
// runs the decode kernel eight columns at a time (same GPU arithmetic)
cudaError_t launch_gemm_q4k_q8k(const void *, int64_t, int64_t, int64_t, const void *, int64_t, int64_t,
                                float *, int64_t, cudaStream_t) {
    return cudaErrorNotSupported;
}

// workspace layout of the prefill path: int8 tiles | mins tiles | float d   (q4k_mul_mat_workspace_bytes adds them up)
size_t gemm_workspace_bytes(int64_t K, int64_t N) {
    const size_t nb = (size_t)(K / QK_K), ntiles = ((size_t)N + BN - 1) / BN;
    return ntiles * BN * (size_t)K + ntiles * nb * MIN_TILE_BYTES + nb * ntiles * BN * 4 + FLAG_BYTES + MAX_CTAS * PARTIAL_BYTES;      // d_x rows are padded to whole column tiles (512-byte bulk copies)
}

// The persistent grid.  More tiles than SMs: one CTA per SM, the units cut evenly (no partial last wave: 448 tiles on 148 SMs
// were four waves for three waves' work).  Fewer tiles than SMs: k CTAs per tile while a part keeps at least 8 superblocks — a
// split costs a parked tile, a flag and a second pipeline fill, which a short K does not earn back (2048 x 2048 x 512, 64 tiles of
// 8 superblocks: 21.7 us whole, 24.4 us cut over 148 CTAs), and with every SM on the tensor pipe the chip runs into its power
// cap, so SMs added to a launch that already fills most of them buy less than their count (4096 x 14336 x 512: 128 tiles whole
// 95 us, cut over 148 CTAs 102 us).  The development / test knob gemm_ksplit = k > 0 forces tiles x k CTAs (capped at the SM
// count): another split, another float association, the same integers.
static int choose_grid(int64_t tiles, int64_t nb, int64_t N, int &halves) {
    const int64_t sms = std::min<int64_t>(sm_count(), (int64_t)MAX_CTAS);
    // Few tiles (TinyLlama's attn_output at 512 tokens: 64 tiles on 148 SMs): 64-column half tiles as the work unit — twice the
    // CTAs, each with half the MMAs and half the scale application per superblock, whole K per CTA, nothing parked.  Only when
    // every tile has two halves with real columns.
    // (a long K is cut between two CTAs per tile instead: 2048 x 5632 x 512 36.3 us that way, 37.4 us as half tiles)
    halves = knobs().gemm_ksplit < 1 && 2 * tiles <= sms && nb < 16 && N > HALF_N && (N % BN == 0 || N % BN > HALF_N) ? 1 : 0;
    if (halves) tiles *= 2;
    const int64_t units = tiles * nb;
    int64_t g;
    if (knobs().gemm_ksplit >= 1) g = std::min(sms, tiles * knobs().gemm_ksplit);
    else if (tiles >= sms) g = sms;
    else g = tiles * std::max<int64_t>(1, std::min(sms / tiles, nb / 8));
    return (int)std::max<int64_t>(1, std::min(g, units));
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda)
static cudaError_t make_weight_map(CUtensorMap *map, const void *w, int64_t M, int64_t nb, int64_t w_row_stride_bytes) {
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
    const cuuint64_t dims[2] = {(cuuint64_t)(nb * BLK_Q4K), (cuuint64_t)M};
    const cuuint64_t strides[1] = {(cuuint64_t)w_row_stride_bytes};
    const cuuint32_t box[2] = {(cuuint32_t)BLK_Q4K, (cuuint32_t)BM}, estr[2] = {1, 1};
    const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(w), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

// F32 activations: quantize into `workspace` (gemm_workspace_bytes), then the tensor-core GEMM.  Two launches.
// x == nullptr: the quantized copy in `workspace` is reused (q4k_mul_mat_reuse) — one launch.
// nmat (1..4) matrices with the same K that read the same activations go out as ONE GEMM launch: the row tiles of all of them
// form one grid (TinyLlama's attn_k / attn_v are 2 row tiles each — 8 CTAs on 148 SMs when launched alone, and a launch costs
// ~20 us whatever its size).  dst[i]: N columns of M[i] floats, column stride dst_col_stride_bytes[i].
cudaError_t launch_gemm_q4k_f32_group(int nmat, const void *const *w, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                                      const float *x, int64_t N, int64_t x_row_stride_bytes, float *const *dst,
                                      const int64_t *dst_col_stride_bytes, void *workspace, cudaStream_t stream, const IntDump *dump) {
    if (nmat < 1 || nmat > 4 || (dump && nmat != 1)) return cudaErrorInvalidValue;
    if (N == 0) return cudaSuccess;
    const int64_t nb = K / QK_K;
    const int64_t ntiles = (N + BN - 1) / BN;             // whole column tiles; columns >= N hold stale bytes, never stored
    int8_t *qs = reinterpret_cast<int8_t *>(workspace);
    uint8_t *bmin = reinterpret_cast<uint8_t *>(qs + ntiles * BN * K);
    float *d = reinterpret_cast<float *>(bmin + ntiles * nb * MIN_TILE_BYTES);
    const int64_t Nld = ntiles * BN;                       // d_x rows padded to whole column tiles: the loader copies 512 bytes per tile and superblock
    int *flags = reinterpret_cast<int *>(reinterpret_cast<char *>(d) + (size_t)nb * Nld * 4);
    GemmParams p;
    int64_t mtiles = 0;
    p.nmat = 0;
    for (int i = 0; i < 4; ++i) { p.w[i] = nullptr; p.M[i] = 0; p.dst[i] = nullptr; p.dst_stride[i] = 0; p.tend[i] = 0; }
    for (int i = 0; i < nmat; ++i) {
        if (M[i] == 0) continue;                          // an empty matrix has no tiles
        const int k = p.nmat++;
        const cudaError_t em = make_weight_map(&p.wmap[k], w[i], M[i], nb, w_row_stride_bytes);
        if (em != cudaSuccess) return em;
        p.w[k] = (const uint8_t *)w[i]; p.M[k] = M[i]; p.dst[k] = dst[i]; p.dst_stride[k] = dst_col_stride_bytes[i] / 4;
        mtiles += (M[i] + BM - 1) / BM;
        p.tend[k] = (int)mtiles;
    }
    if (p.nmat == 0) return cudaSuccess;
    if (ntiles * mtiles * nb >= ((int64_t)1 << 31)) return cudaErrorNotSupported;
    int halves = 0;
    const int grid_ctas = choose_grid(ntiles * mtiles, nb, N, halves);
    const int64_t total = nb * N;
    cudaLaunchAttribute pdl_attr[1];
    pdl_attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pdl_attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaError_t e;
    if (!x) {
        // the workspace already holds the quantized copy of these activations (the mul_mat in front of this one read the same
        // src1: q/k/v, gate/up): only the per-CTA flags have to be reset
        if ((e = cudaMemsetAsync(flags, 0, FLAG_BYTES, stream)) != cudaSuccess) return e;
    } else {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)((total + 7) / 8));
        cfg.blockDim = dim3(256);
        cfg.stream = stream;
        cfg.attrs = pdl_attr;
        cfg.numAttrs = knobs().pdl ? 1 : 0;
        int *qflags = flags;
        e = cudaLaunchKernelEx(&cfg, quantize_q8_planar_kernel, x, qs, d, bmin, nb, total, x_row_stride_bytes, Nld, qflags);
        count_launch();
        if (e != cudaSuccess) return e;
    }
    p.w_stride = w_row_stride_bytes; p.nb = (int)nb;
    p.xq = qs; p.xd = d; p.bmin = bmin; p.N = N; p.Nld = Nld;
    p.ntn = (int)ntiles; p.halves = halves; p.units = ntiles * mtiles * nb * (halves ? 2 : 1); p.flags = flags;
    p.partial = reinterpret_cast<float *>(reinterpret_cast<char *>(flags) + FLAG_BYTES);
    p.dump_sumi = dump ? dump->sumi : nullptr; p.dump_summ = dump ? dump->summ : nullptr;
    auto kern = dump ? gemm_q4k_tc_kernel<true> : gemm_q4k_tc_kernel<false>;
    constexpr size_t smem = SMEM_TOTAL;
    if ((e = ensure_max_dynamic_smem((const void *)kern, (int)smem)) != cudaSuccess) return e;
    const dim3 grid((unsigned)grid_ctas);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cfg.attrs = pdl_attr;
    cfg.numAttrs = knobs().pdl ? 1 : 0;
    e = cudaLaunchKernelEx(&cfg, kern, p);
    count_launch();
    return e;
}

cudaError_t launch_gemm_q4k_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                const float *x, int64_t N, int64_t x_row_stride_bytes,
                                float *dst, int64_t dst_col_stride_bytes, void *workspace, cudaStream_t stream,
                                const IntDump *dump) {
    const void *wv[1] = {w};
    float *dv[1] = {dst};
    return launch_gemm_q4k_f32_group(1, wv, &M, K, w_row_stride_bytes, x, N, x_row_stride_bytes, dv, &dst_col_stride_bytes, workspace, stream, dump);
}

}  // namespace q4k

#ifdef Q4K_GEMM_TRACE
extern "C" int q4k_debug_gemm_trace(long long *host) {
    return (int)cudaMemcpyFromSymbol(host, q4k::g_gemm_trace, sizeof(long long) * 16 * 256);
}
#endif
