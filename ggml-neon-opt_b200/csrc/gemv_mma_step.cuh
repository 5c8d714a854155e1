// gemv_mma_step.cuh — the N == 1 superblock step of the tensor-pipe decode kernels ("columns are sub-blocks", see gemv_q4k_mma.cu),
// shared by gemv_q4k_mma.cu and gemv_q4k_n1s.cu so that both produce the same bits for a row.
#pragma once

#include "gemv_stage.cuh"

namespace q4k {

namespace {

constexpr int MROWS = 16;                 // rows per CTA tile
constexpr int KW = 8;                     // consumer warps = superblocks per round

__device__ __forceinline__ void imma_16832(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint2 lds64(const uint8_t *p) { return *reinterpret_cast<const uint2 *>(p); }
__device__ __forceinline__ void named_bar_sync(int id, int threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory"); }

template <int N> struct MmaCfg {
    // shared-memory row pitch of a slot: N == 1 reads 16-byte pieces (quarter-warp = 2 rows x 4 chunks: pitch = 4 words
    // mod 32), N >= 2 reads 8-byte pieces (half-warp = 4 rows x 4 pieces: pitch = 8 words mod 32)
    static constexpr int PITCH = N == 1 ? ROW_BYTES + 16 : ROW_BYTES + 32;
    static constexpr int SLOT = MROWS * PITCH;
    static constexpr int NC = N == 1 ? 1 : 2;            // result columns held per lane and row
};

// ---- one superblock column of a slot, N == 1 ---------------------------------------------------------------------
// acc[0] / accm[0]: row r, acc[1] / accm[1]: row r + 8 (partial over this lane's chunk)
// dump (DUMP instantiations only): the integers of this (tile, superblock) go out as the kernel computed them
struct DumpAt {
    int32_t *p, *sumi, *summ;
    int64_t row0, M;          // first row of the tile, rows of the matrix
    int kb, nb;
};

template <bool DUMP>
__device__ __forceinline__ void sb_step_n1(const uint8_t *slot_sb, const uint8_t *y0, const uint8_t *y1, const uint8_t *yblk, float yd,
                                           int lane, float (&acc)[2], float (&accm)[2], const DumpAt &da) {
    constexpr int PITCH = MmaCfg<1>::PITCH;
    const int r = lane >> 2, c = lane & 3;
    const uint8_t *p0 = slot_sb + r * PITCH, *p1 = p0 + 8 * PITCH;
    const uint4 h0 = lds128(p0), h1 = lds128(p1);
    const uint4 wa = lds128(p0 + 16 + 32 * c), wb = lds128(p0 + 32 + 32 * c);
    const uint4 va = lds128(p1 + 16 + 32 * c), vb = lds128(p1 + 32 + 32 * c);
    const uint4 ya = lds128(y0), yb = lds128(y0 + 16), za = lds128(y1), zb = lds128(y1 + 16);
    const uint32_t w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
    const uint32_t v[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
    const uint32_t b0[8] = {ya.x, ya.y, ya.z, ya.w, yb.x, yb.y, yb.z, yb.w};
    const uint32_t b1[8] = {za.x, za.y, za.z, za.w, zb.x, zb.y, zb.z, zb.w};
    int ce[4] = {0, 0, 0, 0}, co[4] = {0, 0, 0, 0};          // two accumulators: halves the dependent IMMA chain
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        if (t & 1) imma_16832(co, w[t] & 0x0f0f0f0fu, v[t] & 0x0f0f0f0fu, w[t] & 0xf0f0f0f0u, v[t] & 0xf0f0f0f0u, b0[t], b1[t]);
        else       imma_16832(ce, w[t] & 0x0f0f0f0fu, v[t] & 0x0f0f0f0fu, w[t] & 0xf0f0f0f0u, v[t] & 0xf0f0f0f0u, b0[t], b1[t]);
    }
    // C: [0] = (row r, sub-block 2c), [1] = (row r, 2c + 1, x16), [2] / [3] = row r + 8
    const uint4 u0 = unpack_scales_mins(h0), u1 = unpack_scales_mins(h1);
    const int sh = 16 * (c & 1);
    const uint32_t sc0 = ((c < 2 ? u0.x : u0.y) >> sh) & 0xffffu, sc1 = ((c < 2 ? u1.x : u1.y) >> sh) & 0xffffu;
    const uint32_t mn0 = ((c < 2 ? u0.z : u0.w) >> sh) & 0xffffu, mn1 = ((c < 2 ? u1.z : u1.w) >> sh) & 0xffffu;
    const int s0 = ((int)((sc0 & 0xffu) << 4) * (ce[0] + co[0]) + (int)(sc0 >> 8) * (ce[1] + co[1])) >> 4;
    const int s1 = ((int)((sc1 & 0xffu) << 4) * (ce[2] + co[2]) + (int)(sc1 >> 8) * (ce[3] + co[3])) >> 4;
    const int psw = *reinterpret_cast<const int *>(yblk + 256 + 4 * c);      // pair sums of sub-blocks 2c, 2c + 1
    const int m0 = __dp2a_lo(psw, (int)mn0, 0), m1 = __dp2a_lo(psw, (int)mn1, 0);
    if constexpr (DUMP) {
        // C[row][j] IS p[j]: lane (r, c) holds sub-blocks 2c (exact) and 2c + 1 (x16) of rows r and r + 8
        const int64_t r0 = da.row0 + r, r1 = r0 + 8;
        if (r0 < da.M) { int32_t *o = da.p + ((r0 * da.nb) + da.kb) * 8 + 2 * c; o[0] = ce[0] + co[0]; o[1] = (ce[1] + co[1]) >> 4; }
        if (r1 < da.M) { int32_t *o = da.p + ((r1 * da.nb) + da.kb) * 8 + 2 * c; o[0] = ce[2] + co[2]; o[1] = (ce[3] + co[3]) >> 4; }
        int t0 = s0, t1 = s1, u0 = m0, u1 = m1;               // the four chunk lanes of a row add up to sumi / summ
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
            t0 += __shfl_xor_sync(0xffffffffu, t0, o); t1 += __shfl_xor_sync(0xffffffffu, t1, o);
            u0 += __shfl_xor_sync(0xffffffffu, u0, o); u1 += __shfl_xor_sync(0xffffffffu, u1, o);
        }
        if (c == 0) {
            if (r0 < da.M) { da.sumi[r0 * da.nb + da.kb] = t0; da.summ[r0 * da.nb + da.kb] = u0; }
            if (r1 < da.M) { da.sumi[r1 * da.nb + da.kb] = t1; da.summ[r1 * da.nb + da.kb] = u1; }
        }
    }
    const float2 dd0 = __half22float2(*reinterpret_cast<const __half2 *>(&h0.x));   // {d, dmin}
    const float2 dd1 = __half22float2(*reinterpret_cast<const __half2 *>(&h1.x));
    acc[0] = __fmaf_rn(__fmul_rn(yd, dd0.x), (float)s0, acc[0]);
    accm[0] = __fmaf_rn(__fmul_rn(-yd, dd0.y), (float)m0, accm[0]);
    acc[1] = __fmaf_rn(__fmul_rn(yd, dd1.x), (float)s1, acc[1]);
    accm[1] = __fmaf_rn(__fmul_rn(-yd, dd1.y), (float)m1, accm[1]);
}

}  // namespace

}  // namespace q4k
