// gemv_q4k_mma.cu — decode GEMV (batch 1..8), integer dot products on the tensor pipe.
// Replaces the vec_dot loop of ggml_compute_forward_mul_mat_one_chunk (L/ggml/src/ggml-cpu/ggml-cpu.c:1115-1203)
// + ggml_vec_dot_q4_K_q8_K (L/ggml/src/ggml-cpu/arch/x86/quants.c:1741-1823) for N <= 8, with the
// quantize_row_q8_K pass (ggml-cpu.c:1267-1307, L/ggml/src/ggml-quants.c:2555-2592) fused in front.
//
// Why a second decode kernel.  gemv_q4k_tma.cu does the 4-bit x 8-bit products with dp4a: ~590 issue slots per
// superblock-lane, which puts a B200 SM at ~75% of its issue rate when HBM runs at peak — the kernel is as much
// issue-bound as HBM-bound, and a launch whose tiles all land at once (a 2048-row projection) pays the whole
// burst after the data has arrived.  Here the products go through mma.sync.m16n8k32 (u8 x s8 -> s32, IMMA.16832):
// ~7 issue slots per superblock-row instead of ~18, for every batch size up to 8 at the same cost.  It is still an
// HBM-bound streaming kernel — the tensor pipe runs at a few percent of its rate; it is used as a wide dot-product
// unit, the matrix is never reshaped.  The integers are exact (same sumi / summ as the reference); the float stage
// is the reference's per-superblock form  acc += (d*dy)*sumi,  accm += (-dmin*dy)*summ  with a fixed, launch-
// independent association (below), within 1e-5 of ggml_vec_dot_q4_K_q8_K; the bit-identical AVX2 order remains
// available from gemv_q4k_tma.cu (Q4K_GEMV_ORDER=exact).
//
// Work split.  A CTA tile is 16 weight rows (the MMA's m).  Eight consumer warps split K: in every round warp w
// takes superblock 8*round + w of those 16 rows, so a 2048-column matrix is one round and a 2048-row matrix is
// 128 tiles x 8 warps.  Warp 8 is the producer: per round it drops the 16 x 1152 contiguous bytes into a shared
// memory slot with 16 bulk copies (UBLKCP) on an mbarrier; slots are recycled through a second mbarrier that the
// consumer warps arrive on.  The slot's row pitch is 1152 + 16 (or + 32) bytes so that the fragment loads of the
// eight row lanes hit distinct banks.
//
// Fragments, N == 1 ("columns are sub-blocks").  Lane (r = lane / 4, c = lane % 4) owns rows r and r + 8 and the
// 32-byte nibble chunk c of the superblock, i.e. sub-blocks 2c (low nibbles) and 2c + 1 (high nibbles).  MMA t
// (t = 0..7) multiplies word t of every chunk: the k index of the instruction is (chunk, nibble half, byte of the
// word); the B operand carries the activation word (sub-block j, word t) in column j only, zeros elsewhere, so
// after the eight MMAs  C[row][j] = dot(sub-block j of the row, activations of sub-block j)  for all eight j at
// once.  The high-nibble operands are (w & 0xf0f0f0f0), 16 x the value: the lane folds  16*sc_lo*C_lo + sc_hi*C_hi
// and shifts right by 4 (exact).  N >= 2 ("columns are batch columns"): one MMA per sub-block, k = the 32 bytes
// of the sub-block, B column n = activations of batch column n.
//
// Float association (all N): every lane accumulates its own sub-block pair over the superblocks its warp owns, in
// order; at the end of the tile the four chunk lanes are summed (xor 1, xor 2), then the eight warps' partials in
// warp order.  Nothing depends on the grid, the ring depth or the batch size a column is computed in.
#include <algorithm>
#include <cstdlib>

#include "gemv_stage.cuh"
#include "gemv_mma_step.cuh"
#include "q4k_kernels.h"

namespace q4k {

namespace {

// (MROWS, KW, imma_16832, MmaCfg, DumpAt and sb_step_n1 — the N == 1 superblock step — live in gemv_mma_step.cuh: the streamed
// single-column kernel gemv_q4k_n1s.cu runs the same step, so that its bits are this kernel's bits)
constexpr int MMA_THREADS = (KW + 2) * 32;  // 8 consumer warps, the producer warp, the reducer warp
constexpr int PBUF = 4;                   // partial-sum buffers between the consumer warps and the reducer warp

// ---- one superblock column of a slot, 2 <= N <= 8 ------------------------------------------------------------------
// acc[2 * i + k]: row r + 8 i, batch column 2c + k (complete sums: the MMA reduces over the whole sub-block)
// ycol: this lane's B column (batch column lane / 4) block, or a zero block when lane / 4 >= N
// ylo / yhi: the blocks of batch columns 2c and 2c + 1 (for the min products and dy), clamped to a valid column
template <int N, bool DUMP>
__device__ __forceinline__ void sb_step_nn(const uint8_t *slot_sb, const uint8_t *ycol, const uint8_t *ylo, const uint8_t *yhi,
                                           float yd_lo, float yd_hi, int lane, float (&acc)[4], float (&accm)[4], const DumpAt &da) {
    constexpr int PITCH = MmaCfg<N>::PITCH;
    const int r = lane >> 2, c = lane & 3;
    const uint8_t *p0 = slot_sb + r * PITCH, *p1 = p0 + 8 * PITCH;
    const uint4 h0 = lds128(p0), h1 = lds128(p1);
    const uint4 u0 = unpack_scales_mins(h0), u1 = unpack_scales_mins(h1);
    int s[4] = {0, 0, 0, 0};
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint2 w = lds64(p0 + 16 + 32 * g + 8 * c), v = lds64(p1 + 16 + 32 * g + 8 * c);
        const uint2 bl = lds64(ycol + 64 * g + 8 * c), bh = lds64(ycol + 64 * g + 32 + 8 * c);
        int cl[4] = {0, 0, 0, 0}, ch[4] = {0, 0, 0, 0};
        imma_16832(cl, w.x & 0x0f0f0f0fu, v.x & 0x0f0f0f0fu, w.y & 0x0f0f0f0fu, v.y & 0x0f0f0f0fu, bl.x, bl.y);
        imma_16832(ch, w.x & 0xf0f0f0f0u, v.x & 0xf0f0f0f0u, w.y & 0xf0f0f0f0u, v.y & 0xf0f0f0f0u, bh.x, bh.y);
        const uint32_t q0 = (g < 2 ? u0.x : u0.y) >> (16 * (g & 1)), q1 = (g < 2 ? u1.x : u1.y) >> (16 * (g & 1));
        const int lo0 = (int)((q0 & 0xffu) << 4), hi0 = (int)((q0 >> 8) & 0xffu);
        const int lo1 = (int)((q1 & 0xffu) << 4), hi1 = (int)((q1 >> 8) & 0xffu);
        s[0] += lo0 * cl[0] + hi0 * ch[0];
        s[1] += lo0 * cl[1] + hi0 * ch[1];
        s[2] += lo1 * cl[2] + hi1 * ch[2];
        s[3] += lo1 * cl[3] + hi1 * ch[3];
        if constexpr (DUMP) {      // C[row][column] of this MMA pair IS p[2g] (and 16 x p[2g + 1]) of that (row, batch column)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int64_t row = da.row0 + r + 8 * (i >> 1);
                const int col = 2 * c + (i & 1);
                if (row < da.M && col < N) {
                    int32_t *o = da.p + (((int64_t)col * da.M + row) * da.nb + da.kb) * 8 + 2 * g;
                    o[0] = cl[i]; o[1] = ch[i] >> 4;
                }
            }
        }
    }
    const float2 dd0 = __half22float2(*reinterpret_cast<const __half2 *>(&h0.x));
    const float2 dd1 = __half22float2(*reinterpret_cast<const __half2 *>(&h1.x));
    const uint4 pl = lds128(ylo + 256), ph = lds128(yhi + 256);       // the eight pair sums of each column
    // summ = sum_j m[j] * ps[j]: four dp2a per (row, column)
    const int ml0 = __dp2a_hi((int)pl.w, (int)u0.w, __dp2a_lo((int)pl.z, (int)u0.w, __dp2a_hi((int)pl.y, (int)u0.z, __dp2a_lo((int)pl.x, (int)u0.z, 0))));
    const int mh0 = __dp2a_hi((int)ph.w, (int)u0.w, __dp2a_lo((int)ph.z, (int)u0.w, __dp2a_hi((int)ph.y, (int)u0.z, __dp2a_lo((int)ph.x, (int)u0.z, 0))));
    const int ml1 = __dp2a_hi((int)pl.w, (int)u1.w, __dp2a_lo((int)pl.z, (int)u1.w, __dp2a_hi((int)pl.y, (int)u1.z, __dp2a_lo((int)pl.x, (int)u1.z, 0))));
    const int mh1 = __dp2a_hi((int)ph.w, (int)u1.w, __dp2a_lo((int)ph.z, (int)u1.w, __dp2a_hi((int)ph.y, (int)u1.z, __dp2a_lo((int)ph.x, (int)u1.z, 0))));
    if constexpr (DUMP) {
        const int si[4] = {s[0] >> 4, s[1] >> 4, s[2] >> 4, s[3] >> 4}, sm[4] = {ml0, mh0, ml1, mh1};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t row = da.row0 + r + 8 * (i >> 1);
            const int col = 2 * c + (i & 1);
            if (row < da.M && col < N) {
                da.sumi[((int64_t)col * da.M + row) * da.nb + da.kb] = si[i];
                da.summ[((int64_t)col * da.M + row) * da.nb + da.kb] = sm[i];
            }
        }
    }
    acc[0] = __fmaf_rn(__fmul_rn(yd_lo, dd0.x), (float)(s[0] >> 4), acc[0]);
    acc[1] = __fmaf_rn(__fmul_rn(yd_hi, dd0.x), (float)(s[1] >> 4), acc[1]);
    acc[2] = __fmaf_rn(__fmul_rn(yd_lo, dd1.x), (float)(s[2] >> 4), acc[2]);
    acc[3] = __fmaf_rn(__fmul_rn(yd_hi, dd1.x), (float)(s[3] >> 4), acc[3]);
    accm[0] = __fmaf_rn(__fmul_rn(-yd_lo, dd0.y), (float)ml0, accm[0]);
    accm[1] = __fmaf_rn(__fmul_rn(-yd_hi, dd0.y), (float)mh0, accm[1]);
    accm[2] = __fmaf_rn(__fmul_rn(-yd_lo, dd1.y), (float)ml1, accm[2]);
    accm[3] = __fmaf_rn(__fmul_rn(-yd_hi, dd1.y), (float)mh1, accm[3]);
}


// ---- cluster helpers (split-K: the CTAs of a cluster share one tile, CTA rank = K part) -------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_remote_f32(uint32_t raddr, float v) { asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(raddr), "f"(v) : "memory"); }
__device__ __forceinline__ void mbar_arrive_remote(uint32_t rbar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(rbar) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {      // observes arrivals from other CTAs
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

constexpr int SEG_ROUNDS = 4;             // rounds per segment: the unit of the float association along K (8192 columns)
constexpr int XD = 2;                     // split-K: exchange buffers (tiles in flight between the parts and the leader)

// MODE 0: K <= 8192 (one segment; the common decode shapes, everything about segments folds away), 1: several segments,
// 2: several segments split over the CTAs of a cluster
// DUMP: parity instantiation — the same code, and the integers it computed are also written out (q4k_mul_mat_dump)
template <int N, int MODE, bool DUMP = false>
__global__ void __launch_bounds__(MMA_THREADS, 2)
gemv_q4k_mma_kernel(const __grid_constant__ TmaGemvParams p) {
    constexpr bool SPLITK = MODE == 2;
    using Cfg = MmaCfg<N>;
    constexpr int ITEMS = (N * MROWS + 31) / 32;          // results per reducer lane
    extern __shared__ __align__(128) uint8_t smem[];
    const int nb = p.nb, S = p.stages, KP = SPLITK ? p.kparts : 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp == 0) { TICK(0); GTICK(10); }

    // K split: segments of SEG_ROUNDS rounds; part `rank` of the cluster owns segments [seg0, seg1)
    const int rounds = (nb + KW - 1) / KW, nseg = MODE == 0 ? 1 : (rounds + SEG_ROUNDS - 1) / SEG_ROUNDS;
    const int rank = KP > 1 ? (int)cluster_ctarank() : 0;
    const int seg0 = nseg * rank / KP, seg1 = nseg * (rank + 1) / KP;
    const int pseg_max = (nseg + KP - 1) / KP;            // segments of the largest part

    size_t off = 0;
    uint8_t *q8_s = smem;                                   off += (size_t)p.nx * N * q8_row_bytes(nb);
    uint8_t *zero_s = smem + off;                           off += Q8_STRIDE;
    float *d_s = reinterpret_cast<float *>(smem + off);     off += ((size_t)p.nx * N * nb * 4 + 127) & ~(size_t)127;
    off = (off + 127) & ~(size_t)127;
    float *part_s = reinterpret_cast<float *>(smem + off);  off += (size_t)PBUF * KW * N * MROWS * 4;
    float *xbuf = reinterpret_cast<float *>(smem + off);    off += KP > 1 ? (size_t)XD * (KP - 1) * pseg_max * N * MROWS * 4 : 0;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + off);   off += (size_t)S * 8;
    uint64_t *empty = reinterpret_cast<uint64_t *>(smem + off);  off += (size_t)S * 8;
    uint64_t *pfull = reinterpret_cast<uint64_t *>(smem + off);  off += (size_t)PBUF * 8;
    uint64_t *pempty = reinterpret_cast<uint64_t *>(smem + off); off += (size_t)PBUF * 8;
    uint64_t *xfull = reinterpret_cast<uint64_t *>(smem + off);  off += (size_t)XD * 8;      // leader: a tile's remote parts have landed
    uint64_t *xcredit = reinterpret_cast<uint64_t *>(smem + off); off += (size_t)XD * 8;    // parts: the leader has consumed the buffer
    off = (off + 127) & ~(size_t)127;
    uint8_t *slots = smem + off;

    // tiles of this cluster (n_groups counts tiles): dealt round-robin so that the grid sweeps memory front to back
    const int64_t ncl = gridDim.x / KP, cl = blockIdx.x / KP;
    const int64_t tstep = p.interleave ? ncl : 1;
    const int64_t t0 = p.interleave ? cl : p.n_groups * cl / ncl;
    const int64_t t1 = p.interleave ? p.n_groups : p.n_groups * (cl + 1) / ncl;
    // tile -> (matrix, tile inside the matrix); mat only ever moves forward
    auto locate = [&](int64_t tile, int &mat) -> int64_t {
        while (mat + 1 < p.nmat && tile >= p.gend[mat]) ++mat;
        return mat ? tile - p.gend[mat - 1] : tile;
    };

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full[s]), 1); mbar_init(smem_u32(&empty[s]), KW); }
        for (int b = 0; b < PBUF; ++b) { mbar_init(smem_u32(&pfull[b]), KW); mbar_init(smem_u32(&pempty[b]), 1); }
        for (int b = 0; b < XD; ++b) { mbar_init(smem_u32(&xfull[b]), KP > 1 ? KP - 1 : 1); mbar_init(smem_u32(&xcredit[b]), 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < Q8_STRIDE / 4) reinterpret_cast<uint32_t *>(zero_s)[threadIdx.x] = 0u;
    __syncthreads();
    if (KP > 1) cluster_sync_all();          // every CTA's barriers exist before anyone arrives on them remotely
    griddep_launch_dependents();
    if (warp == 0) TICK(1);

    if (warp == KW) {
        // ---- producer warp: weights never depend on the previous kernel in the stream, so it never waits for it ----
        int s = 0;
        uint32_t eparity = 1;                // parity of the phase BEFORE the first one: the first S waits fall through
        int mat = 0;
        const int rd0 = seg0 * SEG_ROUNDS, rd1 = min(seg1 * SEG_ROUNDS, rounds);
        for (int64_t tile = t0; tile < t1; tile += tstep) {
            const int64_t lt = locate(tile, mat);
            int64_t row = lt * MROWS + (lane & 15);
            if (row >= p.Mm[mat]) row = p.Mm[mat] - 1;          // ragged last tile: stay in bounds, result is not stored
            const uint8_t *src = p.wm[mat] + row * p.w_stride;
            for (int rd = rd0; rd < rd1; ++rd) {
                mbar_wait(smem_u32(&empty[s]), eparity);
                const uint32_t bytes = (uint32_t)min(KW, nb - rd * KW) * BLK_Q4K;
                const uint32_t bar = smem_u32(&full[s]);
                if (lane == 0) mbar_expect_tx(bar, bytes * MROWS);
                __syncwarp();
                if (lane < MROWS) {
                    fence_proxy_async();
                    bulk_g2s(smem_u32(slots + (size_t)s * Cfg::SLOT + lane * Cfg::PITCH), src + (int64_t)rd * ROW_BYTES, bytes, bar);
                }
                if (tile == t0 && rd == rd0) TICK(9);
                if (++s == S) { s = 0; eparity ^= 1u; }
            }
        }
        return;
    }

    if (warp == KW + 1) {
        // ---- reducer warp.  Per segment: the eight warps' partials in warp order.  Per tile: the segment sums in K order —
        // its own in registers; with split-K the other parts' segment sums arrive in the leader's exchange buffer (DSMEM)
        // and are added there in the same order, so the result does not depend on how K was split.
        int mat = 0, b = 0, xs = 0;
        uint32_t fparity = 0, xparity = 0, cparity = 1;
        uint32_t ll_ep = 0;
        if (p.ll_world > 0) { griddep_wait(); ll_ep = ll_epoch(p); if (blockIdx.x == 0) ll_presence(p, ll_ep, lane); }
        for (int64_t tile = t0; tile < t1; tile += tstep) {
            const int64_t lt = locate(tile, mat);
            const int64_t m_rows = p.Mm[mat];
            float tot[ITEMS];
            if (rank > 0) mbar_wait_cluster(smem_u32(&xcredit[xs]), cparity);     // the leader is done with this exchange buffer
            for (int seg = seg0; seg < seg1; ++seg) {
                mbar_wait(smem_u32(&pfull[b]), fparity);
#pragma unroll
                for (int it = 0; it < ITEMS; ++it) {
                    const int i = lane + 32 * it;
                    if (i < N * MROWS) {
                        const float *pp = part_s + (size_t)b * KW * N * MROWS + i;
                        float a = pp[0];
#pragma unroll
                        for (int w = 1; w < KW; ++w) a = __fadd_rn(a, pp[(size_t)w * N * MROWS]);
                        if (rank == 0) tot[it] = seg == 0 ? a : __fadd_rn(tot[it], a);
                        else st_remote_f32(mapa_shared(smem_u32(xbuf + (((size_t)xs * (KP - 1) + rank - 1) * pseg_max + (seg - seg0)) * N * MROWS + i), 0), a);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&pempty[b]));
                if (++b == PBUF) { b = 0; fparity ^= 1u; }
            }
            if (tile == t0) TICK(21);
            if (rank > 0) {
                asm volatile("fence.acq_rel.cluster;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive_remote(mapa_shared(smem_u32(&xfull[xs]), 0));
            } else {
                if (KP > 1) {
                    mbar_wait_cluster(smem_u32(&xfull[xs]), xparity);
                    if (tile == t0) TICK(22);
                    for (int g = 1; g < KP; ++g) {
                        const int gs = nseg * (g + 1) / KP - nseg * g / KP;
                        for (int sl = 0; sl < gs; ++sl) {
#pragma unroll
                            for (int it = 0; it < ITEMS; ++it) {
                                const int i = lane + 32 * it;
                                if (i < N * MROWS) tot[it] = __fadd_rn(tot[it], xbuf[(((size_t)xs * (KP - 1) + g - 1) * pseg_max + sl) * N * MROWS + i]);
                            }
                        }
                    }
                }
#pragma unroll
                for (int it = 0; it < ITEMS; ++it) {
                    const int i = lane + 32 * it;
                    const int col = i / MROWS, rr = i % MROWS;
                    const int64_t row = lt * MROWS + rr;
                    if (i < N * MROWS && row < m_rows) {
                        if (p.ll_world > 0) {
                            ll_store(p, ll_ep, col, p.dst_row0 + row, tot[it]);      // {value, epoch} pairs into every rank's LL buffer
                        } else if (p.nmat == 1) {
                            // row-sharded multi-GPU: the epilogue IS the all-gather — one store per peer over NVLink
                            const int64_t o = (int64_t)col * p.dst_stride + p.dst_row0 + row;
                            p.dst[0][o] = tot[it];
                            for (int g = 1; g < p.n_dst; ++g) p.dst[g][o] = tot[it];
                        } else {
                            p.dstm[mat][(int64_t)col * m_rows + row] = tot[it];     // fused launches write dense [N][M_i] results
                        }
                    }
                }
                if (tile == t0) TICK(23);
                if (KP > 1 && tile + XD * tstep < t1) {      // hand the buffer back (only if a part will use it again: it may have exited)
                    __syncwarp();
                    if (lane > 0 && lane < KP) mbar_arrive_remote(mapa_shared(smem_u32(&xcredit[xs]), lane));
                }
            }
            if (++xs == XD) { xs = 0; xparity ^= 1u; cparity ^= 1u; }
        }
        if (p.sync_world > 0) peer_sync_tail(p, lane);        // row-sharded multi-GPU: the cross-GPU step, fused
        if (p.ll_world > 0 && rank == 0) ll_tail(p, N, ll_ep, cl, ncl, lane, 32);     // the leaders of the clusters share the gather
        return;
    }

    // ---- consumer warps: everything that does not depend on the activations happens before the dependency wait
    // (parameter loads through the constant cache are ~50 cycles each and chain; the compiler would sink them) ----
    const int r = lane >> 2, c = lane & 3;
    const int kb0 = seg0 * SEG_ROUNDS * KW, kb1 = min(seg1 * SEG_ROUNDS * KW, nb);       // this part's superblocks
    StageF32<N> stage;
    stage.prepare(p, KW * 32, kb0, kb1 - kb0);
    touch_params(p);
    int mat = 0;
    locate(t0, mat);
    int abase = p.xslot[mat] * N;
    const int nmat = p.nmat;
    asm volatile("" : "+r"(mat), "+r"(abase));                              // materialised here, not after the wait
    asm volatile("" ::"r"(nmat));
    griddep_wait();
    if (warp == 0) { TICK(2); GTICK(11); }
    if (p.xq) stage_q8<N>(p, q8_s, d_s, KW * 32, kb0, kb1 - kb0);
    else stage.run(q8_s, d_s);
    if (warp == 0) TICK(3);
    named_bar_sync(1, KW * 32);
    if (warp == 0) TICK(4);

    int s = 0, buf = 0;
    uint32_t parity = 0, pparity = 1;        // pparity: phase before the first one, the first PBUF waits fall through
    for (int64_t tile = t0; tile < t1; tile += tstep) {
        DumpAt da{};
        if constexpr (DUMP) {
            int dmat = 0;
            da.p = p.dump_p; da.sumi = p.dump_sumi; da.summ = p.dump_summ;
            da.row0 = locate(tile, dmat) * MROWS; da.M = p.Mm[0]; da.nb = nb; da.kb = 0;
        }
        for (int seg = seg0; seg < seg1; ++seg) {
            float acc[2 * Cfg::NC], accm[2 * Cfg::NC];
#pragma unroll
            for (int i = 0; i < 2 * Cfg::NC; ++i) { acc[i] = 0.0f; accm[i] = 0.0f; }

            const int rde = min((seg + 1) * SEG_ROUNDS, rounds);
            for (int rd = seg * SEG_ROUNDS; rd < rde; ++rd) {
                if (warp == 0 && tile == t0 && rd == seg0 * SEG_ROUNDS) TICK(14);
                mbar_wait(smem_u32(&full[s]), parity);
                if (warp == 0 && tile == t0 && rd == seg0 * SEG_ROUNDS) TICK(5);
                const int kb = rd * KW + warp;
                if (kb < nb) {
                    if constexpr (DUMP) da.kb = kb;
                    const uint8_t *slot_sb = slots + (size_t)s * Cfg::SLOT + warp * BLK_Q4K;
                    if constexpr (N == 1) {
                        const uint8_t *yblk = q8_s + q8_off(abase, kb, nb);
                        const int j = r;                                     // this lane's B column = sub-block j
                        const uint8_t *y0 = (j == 2 * c) ? yblk + 32 * j : zero_s;
                        const uint8_t *y1 = (j == 2 * c + 1) ? yblk + 32 * j : zero_s;
                        sb_step_n1<DUMP>(slot_sb, y0, y1, yblk, d_s[abase * nb + kb], lane, acc, accm, da);
                    } else {
                        const int ncol = r;                                  // this lane's B column = batch column
                        const uint8_t *ycol = ncol < N ? q8_s + q8_off(abase + ncol, kb, nb) : zero_s;
                        const int nlo = min(2 * c, N - 1), nhi = min(2 * c + 1, N - 1);
                        sb_step_nn<N, DUMP>(slot_sb, ycol, q8_s + q8_off(abase + nlo, kb, nb),
                                            q8_s + q8_off(abase + nhi, kb, nb),
                                            d_s[(abase + nlo) * nb + kb], d_s[(abase + nhi) * nb + kb], lane, acc, accm, da);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&empty[s]));
                if (++s == S) { s = 0; parity ^= 1u; }
            }

            if (warp == 0 && tile == t0 && seg == seg0) TICK(6);
            // ---- segment epilogue: chunk lanes here, the eight warps in order in the reducer warp ----
            mbar_wait(smem_u32(&pempty[buf]), pparity);
            float *part = part_s + (size_t)buf * KW * N * MROWS + (size_t)warp * N * MROWS;
            if constexpr (N == 1) {
                float v0 = __fadd_rn(acc[0], accm[0]), v1 = __fadd_rn(acc[1], accm[1]);
                v0 = __fadd_rn(v0, __shfl_xor_sync(0xffffffffu, v0, 1));
                v1 = __fadd_rn(v1, __shfl_xor_sync(0xffffffffu, v1, 1));
                v0 = __fadd_rn(v0, __shfl_xor_sync(0xffffffffu, v0, 2));
                v1 = __fadd_rn(v1, __shfl_xor_sync(0xffffffffu, v1, 2));
                if (c == 0) { part[r] = v0; part[r + 8] = v1; }
            } else {
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const int col = 2 * c + k;
                    if (col < N) {
                        part[col * MROWS + r] = __fadd_rn(acc[k], accm[k]);
                        part[col * MROWS + r + 8] = __fadd_rn(acc[2 + k], accm[2 + k]);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&pfull[buf]));
            if (++buf == PBUF) { buf = 0; pparity ^= 1u; }
        }
        if (tile + tstep < t1) {                                           // the next tile's matrix
            locate(tile + tstep, mat);
            abase = p.xslot[mat] * N;
        }
        if (warp == 0 && tile + tstep >= t1) { TICK(7); GTICK(12); }
    }
}

// ---------------------------------------------------------------------------------------------------- launch
constexpr size_t MMA_SMEM_MAX = 232448;      // 227 KB per CTA on sm_100

size_t mma_fixed_smem(int nrows, int nb, int N, int S, int KP) {
    const int rounds = (nb + KW - 1) / KW, nseg = (rounds + SEG_ROUNDS - 1) / SEG_ROUNDS, pseg_max = (nseg + KP - 1) / KP;
    size_t off = (size_t)nrows * q8_row_bytes(nb) + Q8_STRIDE;
    off += ((size_t)nrows * nb * 4 + 127) & ~(size_t)127;
    off = (off + 127) & ~(size_t)127;
    off += (size_t)PBUF * KW * N * MROWS * 4 + (size_t)S * 16 + (size_t)PBUF * 16 + (size_t)XD * 16;
    if (KP > 1) off += (size_t)XD * (KP - 1) * pseg_max * N * MROWS * 4;
    return ((off + 127) & ~(size_t)127) + 128;
}

// How many CTAs share a tile along K.  Split when one CTA per tile would leave SMs idle or would make every CTA quantize
// a long activation vector by itself; the float result does not depend on the choice (segment order, see the kernel).
int choose_kparts(int64_t tiles, int nb, int nrows /* staged activation rows */, int slots) {
    const int forced = knobs().mma_kparts;
    const int rounds = (nb + KW - 1) / KW, nseg = (rounds + SEG_ROUNDS - 1) / SEG_ROUNDS;
    if (forced > 0) return forced <= nseg ? forced : 1;
    // Rough per-launch model in microseconds, fitted to profiles/r1_launch_chain_splitk.txt: every CTA quantizes the
    // activations of its K part (0.45 us per 8 blocks), then walks its rounds (0.3 us each); two CTAs on one SM slow each
    // other down.  Splitting only pays while all CTAs of the launch are resident at once (one wave).
    int best = 1;
    double best_t = 1e30;
    for (int kp = 1; kp <= 4; kp *= 2) {
        if (kp > nseg) break;
        // beyond one wave only a 2-way split for batches (>= 4 staged rows): halving the activations a CTA stages is worth more
        // than the second wave there (measured: 4096 x 14336 N = 8 30.1 -> 25.8 us, 8192 x 28672 N = 4 62 -> 57 us; no gain
        // at N <= 2 — profiles/r1_batch_kparts.txt)
        // ... and for one column of a long fused-quantizer K (56..64 superblocks: every CTA would quantize all of them):
        // 4096 x 14336 N = 1 12.7 -> 11.8 us (profiles/r1_n1_kparts.txt)
        if (kp > 1 && tiles * kp > slots && !(kp == 2 && (nrows >= 4 || (nrows * nb >= 56 && nb <= 64)))) break;
        const bool two = std::min<int64_t>(tiles * kp, slots) > slots / 2;
        const int prounds = std::min(rounds, (nseg + kp - 1) / kp * SEG_ROUNDS);
        const int blocks = std::min(nb, prounds * KW) * nrows;
        const double stage = 0.45 * ((blocks + KW - 1) / KW) * (two ? 1.6 : 1.0);
        const double waves = (double)((tiles * kp + slots - 1) / slots);
        const double t = stage + waves * prounds * 0.3 * (two ? 1.3 : 1.0) + (kp > 1 ? 0.5 : 0.0);
        if (t < best_t - 1e-9) { best_t = t; best = kp; }
    }
    return best;
}

template <int N>
cudaError_t launch_mma_n(TmaGemvParams p, bool pdl, cudaStream_t stream) {
    const size_t slot = MmaCfg<N>::SLOT;
    const bool dump = p.dump_p != nullptr;
    const int KP = dump ? 1 : choose_kparts(p.n_groups, p.nb, N * p.nx, sm_count() * 2);   // a split part runs the same per-superblock code
    p.kparts = KP;
    const int mode = KP > 1 ? 2 : ((p.nb + KW - 1) / KW <= SEG_ROUNDS ? 0 : 1);
    auto kern = mode == 2 ? gemv_q4k_mma_kernel<N, 2> : (mode == 1 ? gemv_q4k_mma_kernel<N, 1> : gemv_q4k_mma_kernel<N, 0>);
    if (dump) kern = mode == 1 ? gemv_q4k_mma_kernel<N, 1, true> : gemv_q4k_mma_kernel<N, 0, true>;
    const size_t fixed = mma_fixed_smem(N * p.nx, p.nb, N, 8, KP);
    // two CTAs per SM with as deep a ring as fits (<= 5 slots: 2 x 5 x 18 KB in flight per SM is past the latency-
    // bandwidth product); one CTA per SM when the staged activations are too large for that
    int per_sm = 2;
    int stages = fixed + 1024 < MMA_SMEM_MAX / 2 ? (int)((MMA_SMEM_MAX / 2 - 1024 - fixed) / slot) : 0;
    if (stages < 2) {
        per_sm = 1;
        stages = fixed < MMA_SMEM_MAX ? (int)((MMA_SMEM_MAX - fixed) / slot) : 0;
        if (stages < 2) return cudaErrorNotSupported;              // activations do not fit: caller tiles or falls back
    }
    if (stages > 5) stages = 5;
    if (knobs().mma_stages > 0 && knobs().mma_stages < stages) stages = knobs().mma_stages;
    if (knobs().mma_per_sm > 0) per_sm = min(per_sm, knobs().mma_per_sm);
    p.stages = stages;
    p.interleave = knobs().mma_contig > 0 ? 0 : 1;
    const size_t smem = fixed + (size_t)stages * slot;
    if (const cudaError_t ea = ensure_max_dynamic_smem((const void *)kern, (int)MMA_SMEM_MAX)) return ea;
    int64_t grid = (int64_t)sm_count() * per_sm;
    if (knobs().grid_balance & 4) grid = balanced_grid(p.n_groups, grid / KP) * KP;      // (measured slower here: off by default)
    if (grid > p.n_groups * KP) grid = p.n_groups * KP;
    grid -= grid % KP;                               // whole clusters
    if (grid < KP) grid = KP;

    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(MMA_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (KP > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = KP; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
    count_launch();
    if (e == cudaSuccess && p.sync_world > 0 && t_sync_req) t_sync_req->taken = true;
    return e;
}

}  // namespace

// One launch of the MMA decode kernel: nmat (1..4) matrices with the same K and N.  Activations: xq (Q8_K rows, nmat == 1
// only) or xf[i] (F32, quantized in the kernel; equal pointers share one staged copy).  nmat == 1 stores at
// dst[0] + dst_row0 with column stride dst_col_stride_bytes into every peer buffer; nmat > 1 stores dense [N][M_i].
// cudaErrorNotSupported when the staged activations do not fit in shared memory.
cudaError_t launch_gemv_q4k_mma(int nmat, const void *const *w, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                                const void *xq, const float *const *xf, int N, int64_t x_row_stride_bytes,
                                float *const *dst, int64_t dst_col_stride_bytes, float *const *peer_dst, int n_peers,
                                int64_t dst_row0, bool pdl, cudaStream_t stream, const IntDump *dump, const LLExchange *ll) {
    if (nmat < 1 || nmat > 4 || N < 1 || N > 8 || n_peers > 8) return cudaErrorInvalidValue;
    if (dump && nmat != 1) return cudaErrorInvalidValue;
    TmaGemvParams p;
    p.w = (const uint8_t *)w[0]; p.M = M[0]; p.w_stride = w_row_stride_bytes; p.nb = (int)(K / QK_K);
    p.xq = (const uint8_t *)xq; p.xf = xf ? xf[0] : nullptr; p.x_stride = x_row_stride_bytes;
    for (int g = 0; g < 8; ++g) p.dst[g] = nullptr;
    if (n_peers > 0) { for (int g = 0; g < n_peers; ++g) p.dst[g] = peer_dst[g]; p.n_dst = n_peers; }
    else { p.dst[0] = dst ? dst[0] : nullptr; p.n_dst = 1; }
    p.dst_row0 = dst_row0;
    p.dst_stride = dst_col_stride_bytes > 0 ? dst_col_stride_bytes / 4 : M[0];      // 0 = dense [N][M]
    p.stages = 0; p.nmat = nmat; p.nx = 0;
    fill_sync(p, nmat == 1 ? n_peers : 0);
    if (ll && nmat != 1) return cudaErrorInvalidValue;
    fill_ll(p, ll);
    if (dump) { p.dump_p = dump->p; p.dump_sumi = dump->sumi; p.dump_summ = dump->summ; }
    for (int i = 0; i < 4; ++i) { p.wm[i] = nullptr; p.Mm[i] = 0; p.gend[i] = 0; p.dstm[i] = nullptr; p.xsrc[i] = nullptr; p.xslot[i] = 0; }
    int64_t tiles = 0;
    for (int i = 0; i < nmat; ++i) {
        p.wm[i] = (const uint8_t *)w[i]; p.Mm[i] = M[i]; p.dstm[i] = dst ? dst[i] : nullptr;
        tiles += (M[i] + MROWS - 1) / MROWS;
        p.gend[i] = tiles;
        const float *xi = xf ? xf[i] : nullptr;
        int slot = -1;
        for (int j = 0; j < p.nx; ++j) if (p.xsrc[j] == xi) slot = j;
        if (slot < 0) { slot = p.nx; p.xsrc[p.nx++] = xi; }
        p.xslot[i] = slot;
    }
    p.n_groups = tiles;
    if (tiles == 0) return cudaSuccess;
    if (tiles > 0x7fffffff) return cudaErrorInvalidValue;
    switch (N) {
        case 1: return launch_mma_n<1>(p, pdl, stream);
        case 2: return launch_mma_n<2>(p, pdl, stream);
        case 3: return launch_mma_n<3>(p, pdl, stream);
        case 4: return launch_mma_n<4>(p, pdl, stream);
        case 5: return launch_mma_n<5>(p, pdl, stream);
        case 6: return launch_mma_n<6>(p, pdl, stream);
        case 7: return launch_mma_n<7>(p, pdl, stream);
        default: return launch_mma_n<8>(p, pdl, stream);
    }
}

}  // namespace q4k
