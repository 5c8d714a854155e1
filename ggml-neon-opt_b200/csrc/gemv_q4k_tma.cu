// gemv_q4k_tma.cu — decode GEMV (batch 1..8), TMA-bulk staged.  dst[n][m] = vec_dot(W row m, Q8_K(x row n)).
// Replaces ggml_compute_forward_mul_mat_one_chunk's vec_dot loop (L/ggml/src/ggml-cpu/ggml-cpu.c:1115-1203)
// + ggml_vec_dot_q4_K_q8_K (L/ggml/src/ggml-cpu/arch/x86/quants.c:1741-1823) for N <= 8, and — when the
// activations are handed over as F32 — the quantize_row_q8_K pass in front of it (ggml-cpu.c:1267-1307,
// L/ggml/src/ggml-quants.c:2555-2592) inside the same launch.
//
// HBM-bound by design; no tensor cores.  Every weight byte crosses HBM once, moved by the copy engine:
//   * each warp owns a private ring of S slots; a slot holds a 4-row x 8-superblock tile of block_q4_K
//     (4 x 1152 contiguous bytes) fetched with cp.async.bulk (UBLKCP) + mbarrier complete_tx.
//   * one lane per superblock ("lane (r, b)": row r of the tile, superblock b).  The lane pulls its 144 bytes
//     out of shared memory into registers with nine conflict-free LDS.128 (144 B block stride = 36 words =
//     4 banks per lane step), the warp hands the slot straight back to the copy engine, and the arithmetic
//     runs from registers while the next tile is in flight.  The 16 B header is unpacked once per superblock.
//   * the Q8_K activations live in shared memory once per CTA: 272 B per block (256 qs + the 8 pair-summed
//     bsums), 68 words = 4 banks per block step, so eight lanes on eight blocks never conflict and the four
//     rows of a tile broadcast.  They are either copied from a Q8_K workspace or quantized in place from F32
//     (same arithmetic as quantize_q8k.cu, bit-exact).
//   * 4-row groups are dealt round-robin to the CTAs (the grid sweeps the matrix front to back); warps of a CTA pull
//     their CTA's groups from a shared-memory counter, so the tail inside an SM is one row group.
//   * programmatic dependent launch: barrier init and the first weight tiles are issued BEFORE
//     griddepcontrol.wait (weights never depend on the previous kernel), only the activations wait.
//
// Float order.  EXACT reproduces the reference's AVX2 accumulation bit for bit: the integer lane sums
// sumi_t (t = 0..7: word t of every 32-byte nibble group) and the four min products of every superblock
// are parked in a per-warp scratch (transposed, conflict-free), then lane (r, t) runs the reference's FMA
// chain over the superblocks in order: acc_t = fma(d, (float)sumi_t, acc_t), acc_m = fma(dmin, (float)prod,
// acc_m); the final horizontal adds follow hsum_float_8 / movehl / movehdup.  Without EXACT every lane folds
// its superblock into one int and lanes are summed at the end (same integers; fp32 within 1e-5).
//
// The high nibbles are never shifted: dp4a.u32.s32 on (w & 0xf0f0f0f0) gives 16 x the high-nibble dot
// product, the low side is multiplied by 16 x scale, and the (exact) >> 4 happens once per lane sum.
#include <cstdlib>

#include "q4k_common.cuh"
#include "q4k_kernels.h"

#include "gemv_stage.cuh"

namespace q4k {

thread_local PeerSyncReq *t_sync_req = nullptr;

namespace {

// ORD: 0 = fast (one int per lane and superblock), 1 = EXACT (the reference's AVX2 chain), 2 = the float association of
// the tensor-pipe decode kernel (gemv_q4k_mma.cu, N == 1): per lane (= superblock class b of the eight of a round) four
// chains over the 32-byte nibble chunks, summed (c0 + c1) + (c2 + c3) at the end of every 32-superblock segment, the
// eight classes in order, the segments in K order — bit-identical to that kernel, so large matrices can take this
// (copy-engine-bound, faster above ~34 MB) kernel without a row's bits depending on the size of its matrix.
// DUMP: parity instantiation — the same code, and the integers it computed are also written out (q4k_mul_mat_dump)
template <int N, int ORD, bool DUMP = false>
__global__ void __launch_bounds__(WARPS * 32, 2)
gemv_q4k_tma_kernel(const __grid_constant__ TmaBoxParams p) {
    constexpr bool EXACT = ORD == 1, MMAORD = ORD == 2;
    static_assert(!MMAORD || N == 1, "the tensor-pipe order is only instantiated for N == 1");
    extern __shared__ __align__(128) uint8_t smem[];
    const int nb = p.nb, S = p.stages;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r = lane >> 3, b = lane & 7;        // int phase: row r of the tile, superblock b; chain phase: (r, t = b)

    size_t off = 0;
    uint8_t *q8_s = smem;                                  off += (size_t)p.nx * N * q8_row_bytes(nb);
    float *d_s = reinterpret_cast<float *>(smem + off);    off += ((size_t)p.nx * N * nb * 4 + 127) & ~(size_t)127;
    off = (off + 127) & ~(size_t)127;
    uint8_t *wbuf = smem + off + (size_t)warp * S * STAGE_BYTES;   off += (size_t)WARPS * S * STAGE_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + off) + warp * S;   off += (size_t)WARPS * S * 8;
    int2 *meta = reinterpret_cast<int2 *>(smem + off) + warp * S;   off += (size_t)WARPS * S * 8;
    int *counter = reinterpret_cast<int *>(smem + off);    off += 16;
    int *scratch = reinterpret_cast<int *>(smem + off) + (EXACT ? warp * N * SCR_WORDS : 0);

    // 4-row groups are dealt round-robin to the CTAs (group = blockIdx + k * gridDim): the grid sweeps the matrix front to
    // back, which HBM likes better than 296 separate streams (measured on the tensor-pipe and the Q6_K kernels: +10-25%)
    const int ksteps = (nb + TILE_SB - 1) / TILE_SB;

    if (threadIdx.x == 0) { counter[0] = 0; counter[1] = 0; }     // [0] row groups handed out, [1] warps that have finished (fused peer sync)
    if (lane == 0) {
        for (int s = 0; s < S; ++s) mbar_init(smem_u32(&bars[s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // ---- producer state (warp-uniform) ----
    int64_t cur_rg = -1;
    int prod_ks = ksteps;       // forces a fetch on first use
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
        const int nsb = min(TILE_SB, nb - prod_ks * TILE_SB);
        const uint32_t bytes = (uint32_t)nsb * BLK_Q4K;
        if (p.use_tma) {                               // one tensor copy per tile: 4 rows x 1152 B (zeros past the row end / past M)
            if (lane == 0) {
                int mat = 0;
                int64_t lg = cur_rg;
                while (mat + 1 < p.nmat && cur_rg >= p.gend[mat]) { lg = cur_rg - p.gend[mat]; ++mat; }
                meta[s] = make_int2((int)cur_rg, prod_ks);
                mbar_expect_tx(bar, STAGE_BYTES);
                fence_proxy_async();                  // the warp's generic reads of this slot are done (callers sync the warp first)
                tma_box_g2s(smem_u32(wbuf + (size_t)s * STAGE_BYTES), &p.wmap[mat], prod_ks * (ROW_BYTES / 8), (int)(lg * TILE_ROWS), bar);
            }
            ++prod_ks;
            return;
        }
        if (lane == 0) {
            meta[s] = make_int2((int)cur_rg, prod_ks);
            mbar_expect_tx(bar, bytes * TILE_ROWS);
        }
        __syncwarp();
        if (lane < TILE_ROWS) {
            int mat = 0;
            int64_t lg = cur_rg;                      // row group inside its matrix
            while (mat + 1 < p.nmat && cur_rg >= p.gend[mat]) { lg = cur_rg - p.gend[mat]; ++mat; }
            int64_t row = lg * TILE_ROWS + lane;
            if (row >= p.Mm[mat]) row = p.Mm[mat] - 1;   // ragged last group: stay in bounds, result is not stored
            fence_proxy_async();                      // the warp's generic reads of this slot are done (after __syncwarp)
            bulk_g2s(smem_u32(wbuf + (size_t)s * STAGE_BYTES + lane * ROW_BYTES),
                     p.wm[mat] + row * p.w_stride + (int64_t)prod_ks * ROW_BYTES, bytes, bar);
        }
        ++prod_ks;
    };

    // weights never depend on the previous kernel in the stream: get them moving before the dependency wait
    for (int s = 0; s < S; ++s) produce(s);
    griddep_launch_dependents();
    StageF32<N> stage;
    stage.prepare(p, blockDim.x);
    touch_params(p);
    griddep_wait();
    const uint32_t ll_ep = p.ll_world > 0 ? ll_epoch(p) : 0u;     // every warp reads it before the barrier below, i.e. before this CTA can arrive
    if (p.ll_world > 0 && blockIdx.x == 0 && warp == 0) ll_presence(p, ll_ep, lane);
    if (p.xq) stage_q8<N>(p, q8_s, d_s);
    else stage.run(q8_s, d_s);
    __syncthreads();

    float acc[N], accm[N];
#pragma unroll
    for (int n = 0; n < N; ++n) { acc[n] = 0.0f; accm[n] = 0.0f; }
    float cacc[4] = {0.0f, 0.0f, 0.0f, 0.0f}, caccm[4] = {0.0f, 0.0f, 0.0f, 0.0f}, tot = 0.0f;   // MMAORD: chunk chains, tile total

    int s = 0;
    uint32_t parity = 0;
    for (;;) {
        mbar_wait(smem_u32(&bars[s]), parity);
        const int2 mt = meta[s];
        if (mt.x < 0) break;
        const int ks = mt.y;
        const int nsb = min(TILE_SB, nb - ks * TILE_SB);
        const bool active = b < nsb;
        int mat = 0, lgrp = mt.x;
        while (mat + 1 < p.nmat && mt.x >= (int)p.gend[mat]) { lgrp = mt.x - (int)p.gend[mat]; ++mat; }
        const int abase = p.xslot[mat] * N;                 // first staged activation row of this matrix's input

        // the lane's superblock: shared memory -> registers, then the slot goes back to the copy engine
        uint4 hdr, q[8];
        {
            const uint8_t *blk = wbuf + (size_t)s * STAGE_BYTES + r * ROW_BYTES + (active ? b : 0) * BLK_Q4K;
            hdr = lds128(blk);
#pragma unroll
            for (int i = 0; i < 8; ++i) q[i] = lds128(blk + 16 + 16 * i);
        }
        __syncwarp();
        produce(s);

        if (active) {
            const int kb = ks * TILE_SB + b;
            const uint4 sm = unpack_scales_mins(hdr);
            const float2 dd = __half22float2(*reinterpret_cast<const __half2 *>(&hdr.x));   // {d, dmin}
            int sum16[N][EXACT ? 8 : 1];
            int dsumi[N];                                   // DUMP only: the superblock's sumi as this kernel's integers give it
#pragma unroll
            for (int n = 0; n < N; ++n) {
                dsumi[n] = 0;
#pragma unroll
                for (int t = 0; t < (EXACT ? 8 : 1); ++t) sum16[n][t] = 0;
            }
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const uint32_t qw[8] = {q[2 * g].x, q[2 * g].y, q[2 * g].z, q[2 * g].w,
                                        q[2 * g + 1].x, q[2 * g + 1].y, q[2 * g + 1].z, q[2 * g + 1].w};
                const uint32_t scw = (g < 2 ? sm.x : sm.y) >> (16 * (g & 1));
                const int sc_lo16 = (int)((scw & 0xffu) << 4), sc_hi = (int)((scw >> 8) & 0xffu);
#pragma unroll
                for (int n = 0; n < N; ++n) {
                    const uint8_t *yb = q8_s + q8_off(abase + n, kb, nb) + 64 * g;
                    const uint4 y0 = lds128(yb), y1 = lds128(yb + 16), y2 = lds128(yb + 32), y3 = lds128(yb + 48);
                    const uint32_t ylo[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
                    const uint32_t yhi[8] = {y2.x, y2.y, y2.z, y2.w, y3.x, y3.y, y3.z, y3.w};
                    if (EXACT) {
                        int dpl = 0, dph = 0;
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            const int dl = __dp4a((int)(qw[t] & 0x0f0f0f0fu), (int)ylo[t], 0);
                            const int dh = dp4a_u8s8(qw[t] & 0xf0f0f0f0u, yhi[t], 0);
                            sum16[n][t] += sc_lo16 * dl + sc_hi * dh;
                            if constexpr (DUMP) { dpl += dl; dph += dh; }
                        }
                        if constexpr (DUMP) {
                            int32_t *o = p.dump_p + ((((int64_t)n * p.Mm[0] + (int64_t)lgrp * TILE_ROWS + r) * nb) + kb) * 8 + 2 * g;
                            if ((int64_t)lgrp * TILE_ROWS + r < p.Mm[0]) { o[0] = dpl; o[1] = dph >> 4; }
                            dsumi[n] += (sc_lo16 * dpl + sc_hi * dph) >> 4;
                        }
                    } else {
                        int pl = 0, ph = 0;
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            pl = __dp4a((int)(qw[t] & 0x0f0f0f0fu), (int)ylo[t], pl);
                            ph = dp4a_u8s8(qw[t] & 0xf0f0f0f0u, yhi[t], ph);
                        }
                        if constexpr (DUMP) {
                            int32_t *o = p.dump_p + ((((int64_t)n * p.Mm[0] + (int64_t)lgrp * TILE_ROWS + r) * nb) + kb) * 8 + 2 * g;
                            if ((int64_t)lgrp * TILE_ROWS + r < p.Mm[0]) { o[0] = pl; o[1] = ph >> 4; }
                            dsumi[n] += (sc_lo16 * pl + sc_hi * ph) >> 4;
                        }
                        if (MMAORD) {
                            const float dall = __fmul_rn(d_s[(abase + n) * nb + kb], dd.x);
                            cacc[g] = __fmaf_rn(dall, (float)((sc_lo16 * pl + sc_hi * ph) >> 4), cacc[g]);
                        } else {
                            sum16[n][0] += sc_lo16 * pl + sc_hi * ph;
                        }
                    }
                }
            }
#pragma unroll
            for (int n = 0; n < N; ++n) {
                const uint4 ps = lds128(q8_s + q8_off(abase + n, kb, nb) + 256);
                const float yd = d_s[(abase + n) * nb + kb];
                const float dall = __fmul_rn(yd, dd.x), dmn = __fmul_rn(-yd, dd.y);
                // _mm_madd_epi16(mins, hadd(bsums)): prod_i = m[2i] * ps[2i] + m[2i+1] * ps[2i+1]
                const int p0 = __dp2a_lo((int)ps.x, (int)sm.z, 0), p1 = __dp2a_hi((int)ps.y, (int)sm.z, 0);
                const int p2 = __dp2a_lo((int)ps.z, (int)sm.w, 0), p3 = __dp2a_hi((int)ps.w, (int)sm.w, 0);
                if constexpr (DUMP) {
                    const int64_t drow = (int64_t)lgrp * TILE_ROWS + r;
                    if (drow < p.Mm[0]) {
                        p.dump_sumi[((int64_t)n * p.Mm[0] + drow) * nb + kb] = dsumi[n];
                        p.dump_summ[((int64_t)n * p.Mm[0] + drow) * nb + kb] = p0 + p1 + p2 + p3;
                    }
                }
                if (EXACT) {
                    int *row = scratch + n * SCR_WORDS + r * SCR_ROW;
#pragma unroll
                    for (int t = 0; t < 8; ++t) row[t * 8 + (t & 4) + b] = sum16[n][t % (EXACT ? 8 : 1)] >> 4;
                    row[68 + b] = p0; row[76 + b] = p1; row[84 + b] = p2; row[92 + b] = p3;
                    row[100 + b] = __float_as_int(dall);
                    row[108 + b] = __float_as_int(dmn);
                } else if (MMAORD) {
                    caccm[0] = __fmaf_rn(dmn, (float)p0, caccm[0]);
                    caccm[1] = __fmaf_rn(dmn, (float)p1, caccm[1]);
                    caccm[2] = __fmaf_rn(dmn, (float)p2, caccm[2]);
                    caccm[3] = __fmaf_rn(dmn, (float)p3, caccm[3]);
                } else {
                    acc[n] = __fmaf_rn(dall, (float)(sum16[n][0] >> 4), acc[n]);
                    accm[n] = __fmaf_rn(dmn, (float)(p0 + p1 + p2 + p3), accm[n]);
                }
            }
        }
        if (EXACT) {
            __syncwarp();
            // chain phase: lane (r, t = b) walks the superblocks of row r in order, as the AVX2 loop does
#pragma unroll
            for (int n = 0; n < N; ++n) {
                const int *row = scratch + n * SCR_WORDS + r * SCR_ROW;
                const int4 s0 = *reinterpret_cast<const int4 *>(row + b * 8 + (b & 4));
                const int4 s1 = *reinterpret_cast<const int4 *>(row + b * 8 + (b & 4) + 4);
                const int4 m0 = *reinterpret_cast<const int4 *>(row + 68 + (b & 3) * 8);
                const int4 m1 = *reinterpret_cast<const int4 *>(row + 68 + (b & 3) * 8 + 4);
                const float4 da0 = *reinterpret_cast<const float4 *>(row + 100), da1 = *reinterpret_cast<const float4 *>(row + 104);
                const float4 dm0 = *reinterpret_cast<const float4 *>(row + 108), dm1 = *reinterpret_cast<const float4 *>(row + 112);
                const int sv[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
                const int mv[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
                const float da[8] = {da0.x, da0.y, da0.z, da0.w, da1.x, da1.y, da1.z, da1.w};
                const float dm[8] = {dm0.x, dm0.y, dm0.z, dm0.w, dm1.x, dm1.y, dm1.z, dm1.w};
#pragma unroll
                for (int bb = 0; bb < 8; ++bb) {
                    if (bb < nsb) {
                        acc[n] = __fmaf_rn(da[bb], (float)sv[bb], acc[n]);
                        accm[n] = __fmaf_rn(dm[bb], (float)mv[bb], accm[n]);
                    }
                }
            }
        }
        if (MMAORD && ((ks & 3) == 3 || ks == ksteps - 1)) {     // end of a 32-superblock segment (gemv_q4k_mma.cu SEG_ROUNDS)
            const float v01 = __fadd_rn(__fadd_rn(cacc[0], caccm[0]), __fadd_rn(cacc[1], caccm[1]));
            const float v23 = __fadd_rn(__fadd_rn(cacc[2], caccm[2]), __fadd_rn(cacc[3], caccm[3]));
            const float v = __fadd_rn(v01, v23);
            float a = __shfl_sync(0xffffffffu, v, lane & ~7);
#pragma unroll
            for (int w = 1; w < 8; ++w) a = __fadd_rn(a, __shfl_sync(0xffffffffu, v, (lane & ~7) + w));
            tot = (ks >> 2) == 0 ? a : __fadd_rn(tot, a);
#pragma unroll
            for (int c = 0; c < 4; ++c) { cacc[c] = 0.0f; caccm[c] = 0.0f; }
        }
        if (ks == ksteps - 1) {                              // row group finished: horizontal sums + store
            const int64_t row = (int64_t)lgrp * TILE_ROWS + r;
#pragma unroll
            for (int n = 0; n < N; ++n) {
                float a = acc[n], m = accm[n];
                if (EXACT) {
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 4));    // hsum_float_8: t and t+4
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 2));    //   (r0+r2), (r1+r3)
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 1));
                    m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 2));    // acc_m + movehl
                    m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 1));    // + movehdup
                    a = __fadd_rn(a, m);
                } else if (MMAORD) {
                    a = tot;
                } else {
                    a = __fadd_rn(a, m);
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 1));
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 2));
                    a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 4));
                }
                if (b == 0 && row < p.Mm[mat]) {
                    if (p.ll_world > 0) {
                        ll_store(p, ll_ep, n, p.dst_row0 + row, a);              // {value, epoch} pairs into every rank's LL buffer
                    } else if (p.nmat == 1) {
                        // row-sharded multi-GPU: the epilogue IS the all-gather — one store per peer over NVLink
                        const int64_t o = (int64_t)n * p.dst_stride + p.dst_row0 + row;
                        p.dst[0][o] = a;
                        for (int g = 1; g < p.n_dst; ++g) p.dst[g][o] = a;
                    } else {
                        p.dstm[mat][(int64_t)n * p.Mm[mat] + row] = a;      // fused launches write dense [N][M_i] results
                    }
                }
                acc[n] = 0.0f; accm[n] = 0.0f;
            }
        }
        if (++s == S) { s = 0; parity ^= 1u; }
    }
    if (p.ll_world > 0 && warp == 0) ll_tail(p, N, ll_ep, blockIdx.x, gridDim.x, lane, 32);
    if (p.sync_world > 0) {          // row-sharded multi-GPU: the cross-GPU step, fused; the CTA's last warp carries it
        __syncwarp();
        int lastw = 0;
        if (lane == 0) {
            __threadfence_block();                                       // this warp's stores before its arrival (CTA scope: cheap)
            lastw = atomicAdd(&counter[1], 1) == WARPS - 1;
            __threadfence_block();
        }
        lastw = __shfl_sync(0xffffffffu, lastw, 0);
        if (lastw) peer_sync_tail(p, lane);
    }
}

}  // namespace

namespace {

constexpr size_t SMEM_MAX = 232448;      // 227 KB per CTA on sm_100

size_t fixed_smem(int nrows, int nb, bool exact, int N) {    // nrows = staged activation rows (N per distinct input)
    size_t off = (size_t)nrows * q8_row_bytes(nb);
    off += ((size_t)nrows * nb * 4 + 127) & ~(size_t)127;
    off = (off + 127) & ~(size_t)127;
    off += 16;                                               // row-group counter
    if (exact) off += (size_t)WARPS * N * SCR_WORDS * 4;
    return off + 128;
}
constexpr size_t PER_STAGE = (size_t)WARPS * (STAGE_BYTES + 16);

template <int N, int ORD>
cudaError_t launch_tma_n(const TmaGemvParams &base, int want_stages, bool pdl, cudaStream_t stream) {
    TmaBoxParams p;
    static_cast<TmaGemvParams &>(p) = base;
    constexpr bool EXACT = ORD == 1;
    auto kern = p.dump_p ? gemv_q4k_tma_kernel<N, ORD, true> : gemv_q4k_tma_kernel<N, ORD>;
    fill_weight_maps(p, ROW_BYTES, TILE_ROWS);
    const size_t fixed = fixed_smem(N * p.nx, p.nb, EXACT, N);
    if (fixed + PER_STAGE > SMEM_MAX) return cudaErrorNotSupported;      // activations do not fit: caller tiles
    // resident CTAs per SM: as many as registers allow (launch bounds: 2), each with a 2-slot ring when it fits
    int stages = want_stages > 0 ? want_stages : 2;
    if (knobs().tma_stages > 0) stages = knobs().tma_stages;
    while (stages > 1 && 2 * (fixed + stages * PER_STAGE + 1024) > SMEM_MAX) --stages;
    if (fixed + stages * PER_STAGE > SMEM_MAX) stages = (int)((SMEM_MAX - fixed) / PER_STAGE);
    p.stages = stages;
    const size_t smem = fixed + stages * PER_STAGE;
    if (const cudaError_t ea = ensure_max_dynamic_smem((const void *)kern, (int)SMEM_MAX)) return ea;
    int per_sm = (2 * (smem + 1024) <= SMEM_MAX) ? 2 : 1;
    if (knobs().tma_per_sm > 0) per_sm = knobs().tma_per_sm;
    int64_t grid = (int64_t)sm_count() * per_sm;
    if (grid > p.n_groups) grid = p.n_groups;        // small matrices: at least one 4-row group per CTA
    if (grid < 1) grid = 1;

    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(WARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
    count_launch();
    if (e == cudaSuccess && p.sync_world > 0 && t_sync_req) t_sync_req->taken = true;
    return e;
}

template <int EXACT>
cudaError_t launch_tma_e(int N, const TmaGemvParams &p, int stages, bool pdl, cudaStream_t st) {
    switch (N) {
        case 1: return launch_tma_n<1, EXACT>(p, stages, pdl, st);
        case 2: return launch_tma_n<2, EXACT>(p, stages, pdl, st);
        case 3: return launch_tma_n<3, EXACT>(p, stages, pdl, st);
        case 4: return launch_tma_n<4, EXACT>(p, stages, pdl, st);
        case 5: return launch_tma_n<5, EXACT>(p, stages, pdl, st);
        case 6: return launch_tma_n<6, EXACT>(p, stages, pdl, st);
        case 7: return launch_tma_n<7, EXACT>(p, stages, pdl, st);
        case 8: return launch_tma_n<8, EXACT>(p, stages, pdl, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace

// xq != nullptr: activations are Q8_K blocks (x_row_stride_bytes between rows); else xf: F32 rows, quantized
// inside the kernel.  pdl: launch with programmatic stream serialization (the kernel waits for its
// predecessor only before it touches the activations).
cudaError_t launch_gemv_q4k_q8k_tma(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                    const void *xq, const float *xf, int N, int64_t x_row_stride_bytes,
                                    float *dst, int64_t dst_col_stride_bytes, bool exact_order, int stages,
                                    bool pdl, cudaStream_t stream, float *const *peer_dst, int n_peers, int64_t dst_row0,
                                    const IntDump *dump, int dump_kernel, const LLExchange *ll) {
    // dump != nullptr (parity): dump_kernel picks the product kernel whose integers are wanted — 0 = whatever the dispatch
    // below picks, 1 = the tensor-pipe kernel, 2 = the bulk-copy kernel in the order `exact_order` says, 3 = the bulk-copy
    // kernel in the tensor-pipe float association (the large-matrix launch, N == 1)
    if (M == 0 || N == 0) return cudaSuccess;
    if (n_peers > 8) return cudaErrorInvalidValue;
    if (dump && dump_kernel == 3 && N != 1) return cudaErrorInvalidValue;
    // Large single-column launches: the lane-per-superblock kernel in the tensor-pipe kernel's float association (ORD 2,
    // identical bits) — it runs at the bulk-copy ceiling where the 16-row tensor-pipe tiles do not (measured on B200:
    // 28672x8192 24.1 vs 26.3 us, 8192x28672 29.3 vs 33.8 us, 8192x8192 10.3 vs 10.7 us; 4096x14336 14.7 vs 12.5 us, so
    // matrices below 34 MiB stay on the tensor-pipe kernel — profiles/r1_big_sweep.txt).
    // (long K with few rows — a 4096 x 28672 shard — stays on the tensor-pipe kernel, whose cluster split-K wins there:
    // 20.6 vs 22.4 us, profiles/r1_shard_ab_2gpu.txt; the threshold 1 = "always" of the tests skips the row rule)
    const int64_t big_bytes = knobs().big_matrix_bytes;
    bool big = N == 1 && !exact_order && stages <= 0 && big_bytes > 0 && M * (K / QK_K) * BLK_Q4K >= big_bytes &&
               (M >= 8192 || big_bytes == 1);
    bool mma = !big && !exact_order && stages <= 0 && !knobs().mma_off;
    if (dump && dump_kernel == 1) { big = false; mma = true; }
    if (dump && dump_kernel == 2) { big = false; mma = false; }
    if (dump && dump_kernel == 3) { big = true; mma = false; }
    if (mma) {      // tolerance order: the tensor-pipe kernel
        const void *wv[1] = {w};
        const int64_t Mv[1] = {M};
        const float *xv[1] = {xf};
        float *dv[1] = {dst};
        const cudaError_t e = launch_gemv_q4k_mma(1, wv, Mv, K, w_row_stride_bytes, xq, xq ? nullptr : xv, N, x_row_stride_bytes, dv,
                                                  dst_col_stride_bytes, peer_dst, n_peers, dst_row0, pdl, stream, dump, ll);
        if (e != cudaErrorNotSupported || dump) return e;
    }
    TmaGemvParams p;
    p.w = (const uint8_t *)w; p.M = M; p.w_stride = w_row_stride_bytes; p.nb = (int)(K / QK_K);
    p.xq = (const uint8_t *)xq; p.xf = xf; p.x_stride = x_row_stride_bytes;
    for (int g = 0; g < 8; ++g) p.dst[g] = nullptr;
    if (n_peers > 0) { for (int g = 0; g < n_peers; ++g) p.dst[g] = peer_dst[g]; p.n_dst = n_peers; }
    else { p.dst[0] = dst; p.n_dst = 1; }
    p.dst_row0 = dst_row0;
    p.dst_stride = dst_col_stride_bytes / 4;
    p.n_groups = (M + TILE_ROWS - 1) / TILE_ROWS;
    p.stages = 0; p.interleave = 0; p.kparts = 1;
    p.nmat = 1; p.nx = 1;
    fill_sync(p, n_peers);
    fill_ll(p, ll);
    if (dump) { p.dump_p = dump->p; p.dump_sumi = dump->sumi; p.dump_summ = dump->summ; }
    for (int i = 0; i < 4; ++i) { p.wm[i] = nullptr; p.Mm[i] = 0; p.gend[i] = 0; p.dstm[i] = nullptr; p.xsrc[i] = nullptr; p.xslot[i] = 0; }
    p.wm[0] = p.w; p.Mm[0] = M; p.gend[0] = p.n_groups; p.xsrc[0] = xf;
    if (big) {
        const cudaError_t e = launch_tma_n<1, 2>(p, stages, pdl, stream);
        if (e != cudaErrorNotSupported || dump) return e;
        const void *wv[1] = {w};                       // activations too long for this kernel's residency: the tensor-pipe kernel
        const int64_t Mv[1] = {M};
        const float *xv[1] = {xf};
        float *dv[1] = {dst};
        return launch_gemv_q4k_mma(1, wv, Mv, K, w_row_stride_bytes, xq, xq ? nullptr : xv, N, x_row_stride_bytes, dv,
                                   dst_col_stride_bytes, peer_dst, n_peers, dst_row0, pdl, stream, nullptr, ll);
    }
    return exact_order ? launch_tma_e<1>(N, p, stages, pdl, stream) : launch_tma_e<0>(N, p, stages, pdl, stream);
}

namespace {
// the tail of an LL exchange alone (a rank that owns no rows of the matrix still gathers everybody else's)
__global__ void __launch_bounds__(32) ll_gather_kernel(const __grid_constant__ TmaGemvParams p, int N) {
    griddep_launch_dependents();
    griddep_wait();
    const uint32_t ep = ll_epoch(p);
    if (blockIdx.x == 0) ll_presence(p, ep, threadIdx.x);
    ll_tail(p, N, ep, blockIdx.x, gridDim.x, threadIdx.x, 32);
}
}  // namespace

cudaError_t launch_ll_gather(const LLExchange &ll, int N, bool pdl, cudaStream_t stream) {
    TmaGemvParams p = {};
    fill_sync(p, 0);
    fill_ll(p, &ll);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)sm_count());
    cfg.blockDim = dim3(32);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, ll_gather_kernel, p, N);
    count_launch();
    return e;
}

// Horizontally fused decode launch: nmat (<= 4) independent matrices with the same K and N in ONE launch; matrices
// that read the same activations (q/k/v, gate/up) share one staged copy.  F32 activations only; results are dense
// [N][M_i].  cudaErrorNotSupported when the staged activations do not fit.
cudaError_t launch_gemv_q4k_group(int nmat, const void *const *w, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                                  const float *const *x, int N, int64_t x_row_stride_bytes, float *const *dst,
                                  bool exact_order, bool pdl, cudaStream_t stream) {
    if (nmat < 1 || nmat > 4) return cudaErrorInvalidValue;
    if (!exact_order && !knobs().mma_off) {
        const cudaError_t e = launch_gemv_q4k_mma(nmat, w, M, K, w_row_stride_bytes, nullptr, x, N, x_row_stride_bytes, dst,
                                                  0, nullptr, 0, 0, pdl, stream);
        if (e != cudaErrorNotSupported) return e;
    }
    TmaGemvParams p;
    p.w = (const uint8_t *)w[0]; p.M = M[0]; p.w_stride = w_row_stride_bytes; p.nb = (int)(K / QK_K);
    p.xq = nullptr; p.xf = x[0]; p.x_stride = x_row_stride_bytes;
    for (int g = 0; g < 8; ++g) p.dst[g] = nullptr;
    p.dst[0] = dst[0]; p.n_dst = 1; p.dst_row0 = 0; p.dst_stride = M[0];
    p.stages = 0; p.interleave = 0; p.kparts = 1; p.nmat = nmat; p.nx = 0;
    fill_sync(p, 0);
    int64_t groups = 0;
    for (int i = 0; i < 4; ++i) { p.wm[i] = nullptr; p.Mm[i] = 0; p.gend[i] = 0; p.dstm[i] = nullptr; p.xsrc[i] = nullptr; p.xslot[i] = 0; }
    for (int i = 0; i < nmat; ++i) {
        p.wm[i] = (const uint8_t *)w[i]; p.Mm[i] = M[i]; p.dstm[i] = dst[i];
        groups += (M[i] + TILE_ROWS - 1) / TILE_ROWS;
        p.gend[i] = groups;
        int slot = -1;
        for (int j = 0; j < p.nx; ++j) if (p.xsrc[j] == x[i]) slot = j;
        if (slot < 0) { slot = p.nx; p.xsrc[p.nx++] = x[i]; }
        p.xslot[i] = slot;
    }
    p.n_groups = groups;
    if (groups == 0 || N == 0) return cudaSuccess;
    if (groups > 0x7fffffff) return cudaErrorInvalidValue;
    return exact_order ? launch_tma_e<1>(N, p, 0, pdl, stream) : launch_tma_e<0>(N, p, 0, pdl, stream);
}

}  // namespace q4k
