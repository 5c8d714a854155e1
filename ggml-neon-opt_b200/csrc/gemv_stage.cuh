// gemv_stage.cuh — pieces shared by the decode kernels (gemv_q4k_tma.cu, gemv_q4k_mma.cu): launch parameters, the
// mbarrier / bulk-copy / griddepcontrol PTX wrappers, and the staging of the activations in shared memory (Q8_K copied
// from a workspace, or quantized in place from F32 with the arithmetic of quantize_q8k.cu).
#pragma once

#include <cuda.h>

#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

constexpr int TILE_ROWS = 4;
constexpr int TILE_SB = 8;
constexpr int ROW_BYTES = TILE_SB * BLK_Q4K;        // 1152
constexpr int STAGE_BYTES = TILE_ROWS * ROW_BYTES;  // 4608
constexpr int Q8_STRIDE = 272;                      // 256 qs + 8 x int16 pair sums
// Staged activations: row (= batch column, per distinct input) `row` is nb blocks of Q8_STRIDE bytes plus a pad that
// makes the row pitch 8 words mod 32.  Without it the pitch of the usual K (nb = 8, 16, 32, 56, 112) is 0 mod 32 banks
// and the B-fragment loads of the tensor-pipe kernel (lane / 4 = batch column) are 4-way bank conflicts: 62% of the
// shared-memory wavefronts of an N = 8 launch (profiles/r1_ncu_gemv_mma_n8_c5.txt).
__host__ __device__ __forceinline__ int q8_row_bytes(int nb) { return nb * Q8_STRIDE + 4 * ((8 - (nb * (Q8_STRIDE / 4)) % 32 + 32) % 32); }
__host__ __device__ __forceinline__ size_t q8_off(int row, int kb, int nb) { return (size_t)row * q8_row_bytes(nb) + (size_t)kb * Q8_STRIDE; }
constexpr int SCR_ROW = 136;                        // scratch words per (column, tile row); 136 % 32 == 8
constexpr int SCR_WORDS = TILE_ROWS * SCR_ROW;      // per column
constexpr int WARPS = 8;

struct TmaGemvParams {
    const uint8_t *w;
    int64_t M, w_stride;
    int nb;
    const uint8_t *xq;      // Q8_K activations, or
    const float *xf;        // F32 activations (quantized inside the kernel) when xq == nullptr
    int64_t x_stride;       // bytes between activation rows (of whichever is set)
    float *dst[8];          // output buffer of every peer GPU that gets a copy of the result (dst[0] = local)
    int n_dst;
    int64_t dst_row0;       // first row of this shard inside the (full) output column
    int64_t dst_stride;
    int64_t n_groups;       // total 4-row groups of the launch
    int stages;
    int interleave;         // gemv_q4k_mma.cu: 1 = tiles dealt round-robin to the CTAs (the grid sweeps memory front to back)
    int kparts;             // gemv_q4k_mma.cu: CTAs of a thread-block cluster that split K of one tile (1, 2 or 4)
    // horizontally fused launch: up to 4 independent matrices with the same K and N (q/k/v, gate/up ...).  Row groups
    // are numbered through all of them; gend[i] = groups of matrices 0..i.  nmat == 1 uses w / M / dst above.
    int nmat, nx;
    const uint8_t *wm[4];
    int64_t Mm[4], gend[4];
    float *dstm[4];
    const float *xsrc[4];   // the nx distinct F32 activation buffers; xslot[i] = which one matrix i reads
    int xslot[4];
    // fused cross-GPU step of the row-sharded path (q4k_mul_mat_peers_sync): sync_world > 0 = the last CTA of the launch
    // publishes this rank's epoch into every peer's flag row and waits for theirs (peer_sync_tail)
    uint32_t *sync_flags[8];
    uint32_t *sync_epoch, *sync_done;
    int sync_world, sync_rank;
    // parity tooling (DUMP instantiations of the product kernels, nmat == 1): the integer stage of every (column n, row m,
    // superblock b) as the kernel computed it — p[((n*M + m)*nb + b)*8 + j] = the sub-block dot products, sumi / summ
    // [(n*M + m)*nb + b] (L/ggml/src/ggml-cpu/quants.c:575-611: aux32 -> sumi, and the mins sum)
    int32_t *dump_p, *dump_sumi, *dump_summ;
    // LL exchange (q4k_kernels.h LLExchange): ll_world > 0 = results go out as {value, epoch} pairs to every rank's LL buffer
    uint2 *ll_peers[8];
    const uint2 *ll_local;
    int64_t ll_half, ll_M, ll_dst_stride;
    uint32_t *ll_cnt;
    float *ll_dst;
    int ll_world, ll_rank;
};

// gemv_q4k_tma.cu's parameters: + the weight matrices as 2-D tensors of 8-byte elements [M][row bytes / 8] (make_row_box_map):
// ONE tensor copy fetches a whole tile (rows x row piece) where the bulk-copy path needs one copy per row.  (Kept out of the
// struct the tensor-pipe decode kernel takes: 520 more parameter bytes cost its 3.7 us launches 4%.)
struct TmaBoxParams : TmaGemvParams {
    CUtensorMap wmap[4];
    int use_tma;            // 1 = wmap[] are valid and the kernel fetches tiles through them
};

template <class P>
inline void fill_ll(P &p, const LLExchange *ll) {
    for (int r = 0; r < 8; ++r) p.ll_peers[r] = nullptr;
    p.ll_local = nullptr; p.ll_half = p.ll_M = p.ll_dst_stride = 0; p.ll_cnt = nullptr; p.ll_dst = nullptr; p.ll_world = 0; p.ll_rank = 0;
    if (!ll) return;
    for (int r = 0; r < ll->world; ++r) p.ll_peers[r] = static_cast<uint2 *>(ll->peers[r]);
    p.ll_local = static_cast<const uint2 *>(ll->local);
    p.ll_half = ll->half_pairs; p.ll_M = ll->M_full; p.ll_dst_stride = ll->dst_stride;
    p.ll_cnt = ll->cnt; p.ll_dst = ll->dst; p.ll_world = ll->world; p.ll_rank = ll->rank;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda): rows of `row_bytes`
// (a multiple of 8) at `w_row_stride_bytes`, box = box_rows x box_bytes.  false = not available: the caller keeps bulk copies.
inline bool make_row_box_map(CUtensorMap *map, const void *w, int64_t M, int64_t row_bytes, int64_t w_row_stride_bytes, int box_bytes,
                             int box_rows) {
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    static bool tried = false;
    if (!tried) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            encode = (encode_fn)fn;
        tried = true;
    }
    if (!encode || M <= 0 || M > 0x7fffffff || (row_bytes & 7) || (box_bytes & 15) || box_bytes / 8 > 256 || box_rows > 256) return false;
    if ((reinterpret_cast<uintptr_t>(w) & 15) || (w_row_stride_bytes & 15)) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)(row_bytes / 8), (cuuint64_t)M};
    const cuuint64_t strides[1] = {(cuuint64_t)w_row_stride_bytes};
    const cuuint32_t box[2] = {(cuuint32_t)(box_bytes / 8), (cuuint32_t)box_rows}, estr[2] = {1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_INT64, 2, const_cast<void *>(w), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// tensor maps for the nmat matrices of a launch (p.wm / p.Mm / p.w_stride / p.nb must be set); leaves use_tma = 0 on any failure
inline void fill_weight_maps(TmaBoxParams &p, int box_bytes, int box_rows) {
    p.use_tma = 0;
    // Measured (profiles/r2_tma_box_*.txt): worth 1-2% on launches that stream tens of MB (28672 x 8192: 24.41 -> 24.08 us), nothing
    // or slightly negative on small ones (the first tensor copy also fetches its descriptor: 5632 x 2048 3.73 -> 3.81 us).
    // tma_boxes: 0 = never, 1 = launches of at least 24 MiB of weights, 2 = always
    int64_t bytes = 0;
    for (int i = 0; i < p.nmat; ++i) bytes += p.Mm[i] * (int64_t)p.nb * BLK_Q4K;
    if (knobs().tma_boxes == 0 || (knobs().tma_boxes == 1 && bytes < ((int64_t)24 << 20))) return;
    for (int i = 0; i < p.nmat; ++i)
        if (!make_row_box_map(&p.wmap[i], p.wm[i], p.Mm[i], (int64_t)p.nb * BLK_Q4K, p.w_stride, box_bytes, box_rows)) return;
    p.use_tma = 1;
}

inline void fill_sync(TmaGemvParams &p, int n_peers) {
    p.dump_p = p.dump_sumi = p.dump_summ = nullptr;
    fill_ll(p, nullptr);
    for (int r = 0; r < 8; ++r) p.sync_flags[r] = nullptr;
    p.sync_epoch = p.sync_done = nullptr; p.sync_world = 0; p.sync_rank = 0;
    if (t_sync_req && !t_sync_req->taken && n_peers > 0) {
        for (int r = 0; r < 8; ++r) p.sync_flags[r] = t_sync_req->flags[r];
        p.sync_epoch = t_sync_req->epoch; p.sync_done = t_sync_req->done;
        p.sync_world = t_sync_req->world; p.sync_rank = t_sync_req->rank;
    }
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
#ifdef Q4K_TEST_WAIT
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
#else
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
#endif
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// TMA tile copy of a 2-D box global -> shared; c0 = first 8-byte element inside the row, c1 = first row.  Elements outside the
// tensor (past the end of a row, rows past M) arrive as zeros and still count towards the barrier's byte count.
__device__ __forceinline__ void tma_box_g2s(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ int dp4a_u8s8(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint4 lds128(const uint8_t *p) { return *reinterpret_cast<const uint4 *>(p); }

// activations already Q8_K in global memory -> the CTA's shared-memory layout.  [b0, b0 + nbw) = the window of
// superblocks this CTA needs (the whole row unless the launch splits K); nbw = 0 means all.
template <int N>
__device__ __forceinline__ void stage_q8(const TmaGemvParams &p, uint8_t *q8_s, float *d_s, int nt = 0, int b0 = 0, int nbw = 0) {
    if (nt == 0) nt = blockDim.x;              // nt: the first nt threads of the CTA take part
    const int nb = p.nb, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = nt >> 5;
    if (nbw == 0) nbw = nb;
    if (warp >= nwarps) return;
    // one warp per block, four blocks in flight per warp (all loads of a round trip first): a word-per-thread loop with
    // an integer division per word spent 30% of an N = 8 launch here (profiles/r1_ncu_gemv_mma_n8_c5.txt)
    constexpr int R = 4;
    const int total = N * nbw;
    for (int e0 = warp; e0 < total; e0 += R * nwarps) {
        uint32_t a[R], c[R], bw[R];
        float dd[R];
        int blk[R], boff[R];
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
            const int e = e0 + rr * nwarps;
            blk[rr] = -1;
            if (e < total) {                                                   // warp-uniform
                const int n = e / nbw, b = b0 + e - n * nbw;
                blk[rr] = n * nb + b;
                boff[rr] = (int)q8_off(n, b, nb);
                const uint8_t *src = p.xq + n * p.x_stride + (int64_t)b * BLK_Q8K;
                const uint32_t *qs = reinterpret_cast<const uint32_t *>(src + Q8K_QS);
                a[rr] = qs[lane];
                c[rr] = qs[32 + lane];
                bw[rr] = lane < 8 ? reinterpret_cast<const uint32_t *>(src + Q8K_BSUMS)[lane] : 0u;   // bsums 2l, 2l + 1
                dd[rr] = lane == 8 ? *reinterpret_cast<const float *>(src) : 0.0f;
            }
        }
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
            if (blk[rr] >= 0) {
                uint8_t *dst = q8_s + boff[rr];
                reinterpret_cast<uint32_t *>(dst)[lane] = a[rr];
                reinterpret_cast<uint32_t *>(dst)[32 + lane] = c[rr];
                if (lane < 8) reinterpret_cast<int16_t *>(dst + 256)[lane] = (int16_t)((int16_t)(bw[rr] & 0xffffu) + (int16_t)(bw[rr] >> 16));   // hadd_epi16
                if (lane == 8) d_s[blk[rr]] = dd[rr];
            }
        }
    }
}

// F32 activations -> Q8_K in shared memory: one warp per 256-block, the arithmetic of quantize_q8k.cu
// (quantize_row_q8_K_ref, L/ggml/src/ggml-quants.c:2555-2592) through q4k::quantize_block_q8_K.
// Two phases so that nothing but the loads and the arithmetic sits behind griddepcontrol.wait: prepare() reads the
// launch parameters (first touch of a kernel's constant bank is a cache miss) and does the index arithmetic for the
// warp's first STAGE_R blocks; run() issues all loads of a round trip before any arithmetic.
// [b0, b0 + nbw): the window of superblocks this CTA needs (split-K launches stage only their part of K).
constexpr int STAGE_R = 3;

template <int N>
struct StageF32 {
    const float *src[STAGE_R];
    int slot[STAGE_R];                   // shared-memory block index of the prepared blocks (d_s), its byte offset in boff
    int boff[STAGE_R];
    int total, nwarps, nb, b0, nbw;
    const float *x0, *x1, *x2, *x3;      // the (up to four) distinct activation buffers; scalars, not an array: stays in registers
    int64_t x_stride;

    // e-th block of this CTA's window -> source pointer and shared-memory block index
    __device__ __forceinline__ const float *block_src(int e, int &blk, int &off) const {
        const int xs = e / (N * nbw), rem = e - xs * N * nbw, n = rem / nbw, b = b0 + rem - n * nbw;
        const float *x = xs == 0 ? x0 : (xs == 1 ? x1 : (xs == 2 ? x2 : x3));
        blk = (xs * N + n) * nb + b;
        off = (int)q8_off(xs * N + n, b, nb);
        return reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + n * x_stride) + b * QK_K;
    }
    __device__ __forceinline__ void prepare(const TmaGemvParams &p, int nt, int win0 = 0, int win = 0) {
        nb = p.nb; b0 = win0; nbw = win ? win : p.nb;
        nwarps = nt >> 5; total = p.nx * N * nbw; x_stride = p.x_stride;
        x0 = p.xsrc[0]; x1 = p.xsrc[1]; x2 = p.xsrc[2]; x3 = p.xsrc[3];
        const int warp = threadIdx.x >> 5;
#pragma unroll
        for (int rr = 0; rr < STAGE_R; ++rr) {
            const int e = warp + rr * nwarps;
            slot[rr] = 0; boff[rr] = 0;
            src[rr] = e < total ? block_src(e, slot[rr], boff[rr]) : nullptr;
        }
    }
    __device__ __forceinline__ void quantize_to(const float4 &a, const float4 &c, int blk, int off, uint8_t *q8_s, float *d_s) const {
        const int lane = threadIdx.x & 31;
        const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
        uint32_t w_lo, w_hi;
        int s_lo, s_hi;
        const float d = quantize_block_q8_K<8>(v, lane, w_lo, w_hi, s_lo, s_hi);     // s_*: the 32-element pair sums
        uint8_t *dst = q8_s + off;
        reinterpret_cast<uint32_t *>(dst)[lane] = w_lo;
        reinterpret_cast<uint32_t *>(dst)[32 + lane] = w_hi;
        if ((lane & 7) == 0) {
            reinterpret_cast<int16_t *>(dst + 256)[lane >> 3] = (int16_t)s_lo;
            reinterpret_cast<int16_t *>(dst + 256)[4 + (lane >> 3)] = (int16_t)s_hi;
        }
        if (lane == 0) d_s[blk] = d;
    }
    __device__ __forceinline__ void run(uint8_t *q8_s, float *d_s) const {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        {   // the prepared blocks: nothing but loads and arithmetic (no index math: integer division by nb is ~100 instructions)
            float4 a[STAGE_R], c[STAGE_R];
#pragma unroll
            for (int rr = 0; rr < STAGE_R; ++rr) {
                if (src[rr]) {                                                        // warp-uniform
                    a[rr] = reinterpret_cast<const float4 *>(src[rr])[lane];          // plain (coherent) loads: with PDL the producer
                    c[rr] = reinterpret_cast<const float4 *>(src[rr])[32 + lane];     // kernel may still have been running at entry
                }
            }
#ifdef Q4K_TIMING
            if (warp == 0) TICK(13);
#endif
#pragma unroll
            for (int rr = 0; rr < STAGE_R; ++rr)
                if (src[rr]) quantize_to(a[rr], c[rr], slot[rr], boff[rr], q8_s, d_s);
        }
        if (total > STAGE_R * nwarps) {                                               // long activations: the rest
#pragma unroll 1
            for (int base = warp + STAGE_R * nwarps; base < total; base += nwarps * STAGE_R) {
                const float *sp[STAGE_R];
                int sl[STAGE_R], so[STAGE_R];
                float4 a[STAGE_R], c[STAGE_R];
#pragma unroll
                for (int rr = 0; rr < STAGE_R; ++rr) {
                    const int e = base + rr * nwarps;
                    sl[rr] = 0; so[rr] = 0;
                    sp[rr] = e < total ? block_src(e, sl[rr], so[rr]) : nullptr;
                    if (sp[rr]) {
                        a[rr] = reinterpret_cast<const float4 *>(sp[rr])[lane];
                        c[rr] = reinterpret_cast<const float4 *>(sp[rr])[32 + lane];
                    }
                }
#pragma unroll
                for (int rr = 0; rr < STAGE_R; ++rr)
                    if (sp[rr]) quantize_to(a[rr], c[rr], sl[rr], so[rr], q8_s, d_s);
            }
        }
    }
};

template <int N>
__device__ __forceinline__ void stage_f32(const TmaGemvParams &p, uint8_t *q8_s, float *d_s, int nt = 0) {
    StageF32<N> st;
    st.prepare(p, nt == 0 ? blockDim.x : nt);
    st.run(q8_s, d_s);
}

// The cross-GPU step fused into the decode launch (what q4k_peer_signal_wait does in a kernel of its own, peer_sync.cu).
// Called by ONE full warp per CTA after all of the CTA's result stores (to every peer, over NVLink) have been issued and
// fenced.  The last CTA of the grid to arrive bumps the device-side epoch, publishes it with a system-scope release store
// into every peer's flag row and spins until every peer's epoch has arrived here: when the kernel completes, every rank's
// rows have landed in this rank's output buffer.  Saves the dependent launch of the separate signal kernel (~2.5 us).
__device__ __forceinline__ void peer_sync_tail(const TmaGemvParams &p, int lane) {
    // One GPU-scope fence per CTA in front of its arrival; the ONE system-scope fence + release is the last CTA's (causality
    // is transitive: stores -> warp / CTA synchronisation -> this fence -> arrival -> the last CTA's fence.sys -> flag).
    // System-scope fences in every CTA (or worse, every lane) wait for the NVLink acknowledgements 296 times over and cost
    // 8-13 us per launch (profiles/r1_shard_timing_fused_2gpu.txt).
    __syncwarp();
    int last = 0;
    if (lane == 0) {
        __threadfence();
        last = atomicAdd(p.sync_done, 1u) == gridDim.x - 1u;
    }
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    uint32_t e = 0;
    if (lane == 0) {
        __threadfence_system();                                // acquire side of the other CTAs' arrivals
        *p.sync_done = 0u;                                     // everybody has arrived: ready for the next launch
        e = *p.sync_epoch + 1u;
        *p.sync_epoch = e;
    }
    __syncwarp();
    e = __shfl_sync(0xffffffffu, e, 0);
    if (lane < p.sync_world && lane != p.sync_rank) {
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.sync_flags[lane] + p.sync_rank), "r"(e) : "memory");
        const uint32_t *mine = p.sync_flags[p.sync_rank] + lane;
        uint32_t seen;
        do {
            asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
        } while ((int32_t)(seen - e) < 0);                     // wrap-safe "seen >= e"
    }
    __syncwarp();
    asm volatile("fence.acq_rel.sys;" ::: "memory");
}

// ---- LL exchange -------------------------------------------------------------------------------------------------------
// epoch of this call: one more than the calls this object has completed (read after griddepcontrol.wait: the previous call's
// last CTA wrote it)
template <class P>
__device__ __forceinline__ uint32_t ll_epoch(const P &p) {
    uint32_t c;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(c) : "l"(p.ll_cnt) : "memory");
    return c + 1u;
}
// one result -> every rank's LL buffer.  The value and its epoch travel in ONE 8-byte store (atomic over NVLink), so the
// receiver needs no fence: a pair whose epoch matches carries the value.
template <class P>
__device__ __forceinline__ void ll_store(const P &p, uint32_t ep, int64_t col, int64_t row_full, float v) {
    const int64_t i = (int64_t)(ep & 1u) * p.ll_half + col * p.ll_M + row_full;
    for (int g = 0; g < p.ll_world; ++g)
        asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p.ll_peers[g] + i), "r"(__float_as_uint(v)), "r"(ep) : "memory");
}
// "this rank has started call ep" -> every rank (ONE warp of the launch, after it has read the epoch, i.e. after this rank's
// previous call has completed): see the back-pressure note at LLExchange
template <class P>
__device__ __forceinline__ void ll_presence(const P &p, uint32_t ep, int lane) {
    if (lane < p.ll_world) {
        uint2 *slot = p.ll_peers[lane] + 2 * p.ll_half + (int64_t)(ep & 1u) * 8 + p.ll_rank;
        asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(slot), "r"(0u), "r"(ep) : "memory");
    }
}
// The tail: `nlanes` threads of ONE warp-sized group per CTA (lane = 0..nlanes-1) take slice `part` of `nparts` of the gathered
// [N][M_full] vector, wait for each pair of this call to arrive in the rank's own LL buffer (whoever computes it: another
// rank over NVLink, or a CTA of this launch) and write it out as a plain float.  The last CTA to finish publishes the call
// counter.  Every CTA of the launch is resident (persistent grid), so nobody waits for a CTA that cannot run.
template <class P>
__device__ __forceinline__ void ll_tail(const P &p, int N, uint32_t ep, int64_t part, int64_t nparts, int lane, int nlanes) {
    const int64_t total = (int64_t)N * p.ll_M;
    const int64_t i0 = total * part / nparts, i1 = total * (part + 1) / nparts;
    const uint2 *src = p.ll_local + (int64_t)(ep & 1u) * p.ll_half;
    if (part == 0 && lane < p.ll_world) {          // every rank has started this call (= has finished reading the call before)
        const uint2 *slot = p.ll_local + 2 * p.ll_half + (int64_t)(ep & 1u) * 8 + lane;
        uint32_t v, e;
        do {
            asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(e) : "l"(slot) : "memory");
        } while (e != ep);
    }
    for (int64_t i = i0 + lane; i < i1; i += nlanes) {
        uint32_t v, e;
        do {
            asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(e) : "l"(src + i) : "memory");
        } while (e != ep);
        const int64_t col = i / p.ll_M, row = i - col * p.ll_M;
        p.ll_dst[col * p.ll_dst_stride + row] = __uint_as_float(v);
    }
    __syncwarp();
    if (lane == 0) {
        const uint32_t old = atomicAdd(p.ll_cnt + 1, 1u);
        if (old == (uint32_t)nparts - 1u) {            // every CTA has read cnt[0] (it read it before it arrived): publish the new count
            p.ll_cnt[1] = 0u;
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p.ll_cnt), "r"(ep) : "memory");
        }
    }
}

// Reads one field of every 64-byte piece of the launch parameters: the first access to a kernel's constant bank misses
// the constant cache (~0.3 us); done before griddepcontrol.wait, the misses overlap the previous kernel.
__device__ __forceinline__ void touch_params(const TmaGemvParams &p) {
    asm volatile("" ::"l"(p.w), "l"(p.x_stride), "l"(p.dst[3]), "l"(p.dst[7]), "l"(p.dst_row0), "l"(p.dst_stride), "r"(p.nmat),
                 "l"(p.wm[3]), "l"(p.Mm[3]), "l"(p.gend[3]), "l"(p.dstm[3]), "l"(p.xsrc[3]), "r"(p.xslot[3]));
}

}  // namespace

}  // namespace q4k
