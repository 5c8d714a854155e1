/*
 * q4k_oracle.c — CPU ORACLE (test infrastructure, see q4k_oracle.h).
 *
 * Plain-C restatement of the reference's Q4_K x Q8_K mul_mat path.  Build with
 *   gcc -O2 -ffp-contract=off -fno-fast-math      (oracle/Makefile)
 * so that every float operation below is one IEEE-754 binary32 rounding, which is what the
 * reference's x86 build of ggml-base does (no FMA contraction: SURVEY.md §3.4).
 *
 * "L/" = /root/reference/external/llama.cpp/
 */
#include "q4k_oracle.h"

#include <math.h>
#include <string.h>

_Static_assert(sizeof(ora_block_q4_K) == 144, "block_q4_K is 144 bytes (L/ggml/src/ggml-common.h:306)");
_Static_assert(sizeof(ora_block_q8_K) == 292, "block_q8_K is 292 bytes (L/ggml/src/ggml-common.h:344)");
_Static_assert(offsetof(ora_block_q4_K, scales) == 4 && offsetof(ora_block_q4_K, qs) == 16, "q4_K offsets");
_Static_assert(offsetof(ora_block_q8_K, qs) == 4 && offsetof(ora_block_q8_K, bsums) == 260, "q8_K offsets");

/* ------------------------------------------------------------------------------------------------
 * scalar helpers
 * ---------------------------------------------------------------------------------------------- */

/* L/ggml/src/ggml-quants.c:444-449 — round-half-even through the 1.5*2^23 magic constant.
 * Stated here through bits so it does not depend on the FP rounding mode being queried. */
int ora_nearest_int(float f) {
    const float shifted = f + 12582912.0f;            /* one rounding: this IS the rounding step */
    uint32_t u;
    memcpy(&u, &shifted, sizeof u);
    return (int)(u & 0x7fffffu) - 0x400000;
}

/* IEEE binary32 -> binary16, round-to-nearest-even.  Same function as the bit trick in
 * L/ggml/src/ggml-impl.h:401-425 (ggml-base is built without F16C); written here as an explicit
 * integer rounding so the two can be cross-checked over all 2^32 inputs (tests). */
uint16_t ora_fp32_to_fp16(float f) {
    uint32_t u;
    memcpy(&u, &f, sizeof u);
    const uint16_t sign = (uint16_t)((u >> 16) & 0x8000u);
    const uint32_t mag  = u & 0x7fffffffu;
    if (mag > 0x7f800000u) return (uint16_t)(sign | 0x7e00u);        /* NaN  */
    if (mag >= 0x47800000u) return (uint16_t)(sign | 0x7c00u);       /* >= 65536 (incl. inf) -> inf */
    if (mag >= 0x38800000u) {                                        /* normal half range: 2^-14 .. */
        const uint32_t rebias = mag - 0x38000000u;                   /* exponent 127 -> 15 */
        uint32_t h = rebias >> 13;
        const uint32_t rem = rebias & 0x1fffu;
        if (rem > 0x1000u || (rem == 0x1000u && (h & 1u))) h += 1;   /* may carry into inf: correct */
        return (uint16_t)(sign | h);
    }
    if (mag < 0x33000000u) return sign;                              /* < 2^-25 -> 0 */
    /* subnormal half: value = mant * 2^(e-150), half ulp = 2^-24 */
    const int e = (int)(mag >> 23);
    const uint32_t mant = (mag & 0x7fffffu) | 0x800000u;
    const int shift = 126 - e;                                       /* 14..24 */
    uint32_t h = mant >> shift;
    const uint32_t rem = mant & ((1u << shift) - 1u);
    const uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (h & 1u))) h += 1;
    return (uint16_t)(sign | h);
}

/* L/ggml/src/ggml-impl.h:378-399 (exact for every input: fp16 is a subset of fp32) */
float ora_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    const uint32_t e = (h >> 10) & 0x1fu;
    const uint32_t m = h & 0x3ffu;
    uint32_t u;
    if (e == 0x1fu) {
        u = sign | 0x7f800000u | (m << 13);
    } else if (e != 0) {
        u = sign | ((e + 112u) << 23) | (m << 13);
    } else if (m == 0) {
        u = sign;
    } else {
        int k = 0;
        uint32_t mm = m;
        while (!(mm & 0x400u)) { mm <<= 1; ++k; }
        u = sign | ((uint32_t)(113 - k) << 23) | ((mm & 0x3ffu) << 13);
    }
    float f;
    memcpy(&f, &u, sizeof f);
    return f;
}

/* L/ggml/src/ggml-quants.c:703-710 (get_scale_min_k4), all eight at once.  Equivalent to the
 * kmask1/2/3 word shuffle of L/ggml/src/ggml-cpu/quants.c:589-594. */
void ora_unpack_scales_mins(const uint8_t q[12], uint8_t sc[8], uint8_t mn[8]) {
    for (int j = 0; j < 4; ++j) {
        sc[j]     = q[j] & 63;
        mn[j]     = q[j + 4] & 63;
        sc[j + 4] = (uint8_t)((q[j + 8] & 0x0f) | ((q[j] >> 6) << 4));
        mn[j + 4] = (uint8_t)((q[j + 8] >> 4) | ((q[j + 4] >> 6) << 4));
    }
}

/* ------------------------------------------------------------------------------------------------
 * activation quantizer — L/ggml/src/ggml-quants.c:2555-2592 (quantize_row_q8_K_ref; every arch's
 * quantize_row_q8_K forwards to it: arch/x86/quants.c:493-495, arch/arm/quants.c:134-136)
 * ---------------------------------------------------------------------------------------------- */
void ora_quantize_row_q8_K(const float *x, void *vy, int64_t k) {
    ora_block_q8_K *y = (ora_block_q8_K *)vy;
    const int64_t nb = k / ORA_QK_K;
    for (int64_t b = 0; b < nb; ++b, x += ORA_QK_K) {
        float amax = 0.0f, signed_max = 0.0f;
        for (int j = 0; j < ORA_QK_K; ++j) {          /* strict '>' : first occurrence wins */
            const float a = fabsf(x[j]);
            if (a > amax) { amax = a; signed_max = x[j]; }
        }
        if (amax == 0.0f) {                           /* :2568-2573 — bsums are NOT written */
            y[b].d = 0.0f;
            memset(y[b].qs, 0, ORA_QK_K);
            continue;
        }
        const float iscale = -127.0f / signed_max;    /* :2576 */
        for (int j = 0; j < ORA_QK_K; ++j) {
            const int v = ora_nearest_int(iscale * x[j]);
            y[b].qs[j] = (int8_t)(v < 127 ? v : 127);
        }
        for (int g = 0; g < ORA_QK_K / 16; ++g) {
            int s = 0;
            for (int t = 0; t < 16; ++t) s += y[b].qs[16 * g + t];
            y[b].bsums[g] = (int16_t)s;
        }
        y[b].d = 1.0f / iscale;                       /* :2589 */
    }
}

/* ------------------------------------------------------------------------------------------------
 * weight quantizer — L/ggml/src/ggml-quants.c:622-701 (make_qkx2_quants) and :1280-1350
 * (quantize_row_q4_K_ref).  Input generator only: it is what the tests use to make Q4_K tensors
 * identical to the reference's; it is not accelerated on the GPU.
 * ---------------------------------------------------------------------------------------------- */
static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* returns the scale, writes -min to *neg_min; L[n] receives the chosen levels */
static float search_scale_min(int n, int nmax, const float *x, const float *w, uint8_t *L, float *neg_min,
                              uint8_t *Ltry, float rmin, float rdelta, int nstep) {
    float lo = x[0], hi = x[0];
    float sw = w[0];
    float swx = sw * x[0];
    for (int i = 1; i < n; ++i) {
        if (x[i] < lo) lo = x[i];
        if (x[i] > hi) hi = x[i];
        sw += w[i];
        swx += w[i] * x[i];
    }
    if (lo > 0) lo = 0;
    if (hi == lo) {
        memset(L, 0, (size_t)n);
        *neg_min = -lo;
        return 0.0f;
    }
    float inv = (float)nmax / (hi - lo);
    float scale = 1 / inv;
    float best = 0;
    for (int i = 0; i < n; ++i) {
        const int l = clampi(ora_nearest_int(inv * (x[i] - lo)), 0, nmax);
        L[i] = (uint8_t)l;
        float e = scale * L[i] + lo - x[i];
        e = e * e;
        best += w[i] * e;
    }
    for (int step = 0; step <= nstep; ++step) {
        inv = (rmin + rdelta * step + nmax) / (hi - lo);
        float sl = 0, sl2 = 0, sxl = 0;
        for (int i = 0; i < n; ++i) {
            const int l = clampi(ora_nearest_int(inv * (x[i] - lo)), 0, nmax);
            Ltry[i] = (uint8_t)l;
            sl  += w[i] * l;
            sl2 += w[i] * l * l;
            sxl += w[i] * l * x[i];
        }
        const float det = sw * sl2 - sl * sl;
        if (!(det > 0)) continue;
        float s_try = (sw * sxl - swx * sl) / det;
        float m_try = (sl2 * swx - sl * sxl) / det;
        if (m_try > 0) { m_try = 0; s_try = sxl / sl2; }
        float err = 0;
        for (int i = 0; i < n; ++i) {
            float e = s_try * Ltry[i] + m_try - x[i];
            e = e * e;
            err += w[i] * e;
        }
        if (err < best) {
            memcpy(L, Ltry, (size_t)n);
            best = err; scale = s_try; lo = m_try;
        }
    }
    *neg_min = -lo;
    return scale;
}

void ora_quantize_row_q4_K(const float *x, void *vy, int64_t k) {
    ora_block_q4_K *y = (ora_block_q4_K *)vy;
    const int64_t nb = k / ORA_QK_K;
    uint8_t L[ORA_QK_K], Ltry[32];
    float w[32], sub_scale[8], sub_min[8];

    for (int64_t b = 0; b < nb; ++b, x += ORA_QK_K) {
        float top_scale = 0, top_min = 0;
        for (int j = 0; j < 8; ++j) {
            const float *xs = x + 32 * j;
            float ss = 0;
            for (int l = 0; l < 32; ++l) ss += xs[l] * xs[l];
            const float rms = sqrtf(ss / 32);
            for (int l = 0; l < 32; ++l) w[l] = rms + fabsf(xs[l]);
            sub_scale[j] = search_scale_min(32, 15, xs, w, L + 32 * j, &sub_min[j], Ltry, -1.0f, 0.1f, 20);
            if (sub_scale[j] > top_scale) top_scale = sub_scale[j];
            if (sub_min[j] > top_min) top_min = sub_min[j];
        }
        const float inv_s = top_scale > 0 ? 63.0f / top_scale : 0.0f;
        const float inv_m = top_min > 0 ? 63.0f / top_min : 0.0f;
        memset(y[b].scales, 0, sizeof y[b].scales);
        for (int j = 0; j < 8; ++j) {
            uint8_t ls = (uint8_t)ora_nearest_int(inv_s * sub_scale[j]);
            uint8_t lm = (uint8_t)ora_nearest_int(inv_m * sub_min[j]);
            if (ls > 63) ls = 63;
            if (lm > 63) lm = 63;
            if (j < 4) {                               /* 6-bit packing, :1317-1324 */
                y[b].scales[j] |= ls;
                y[b].scales[j + 4] |= lm;
            } else {
                y[b].scales[j + 4] = (uint8_t)((ls & 0x0f) | ((lm & 0x0f) << 4));
                y[b].scales[j - 4] |= (uint8_t)((ls >> 4) << 6);
                y[b].scales[j]     |= (uint8_t)((lm >> 4) << 6);
            }
        }
        y[b].d    = ora_fp32_to_fp16(top_scale / 63.0f);
        y[b].dmin = ora_fp32_to_fp16(top_min / 63.0f);

        uint8_t sc[8], mn[8];
        ora_unpack_scales_mins(y[b].scales, sc, mn);
        const float fd = ora_fp16_to_fp32(y[b].d), fm = ora_fp16_to_fp32(y[b].dmin);
        for (int j = 0; j < 8; ++j) {                  /* re-quantize against the rounded scales, :1329-1339 */
            const float d = fd * sc[j];
            if (d == 0.0f) continue;                   /* keeps the levels found by the search */
            const float dm = fm * mn[j];
            for (int l = 0; l < 32; ++l) {
                L[32 * j + l] = (uint8_t)clampi(ora_nearest_int((x[32 * j + l] + dm) / d), 0, 15);
            }
        }
        for (int g = 0; g < 4; ++g)                    /* nibble packing, :1341-1346 */
            for (int l = 0; l < 32; ++l)
                y[b].qs[32 * g + l] = (uint8_t)(L[64 * g + l] | (L[64 * g + 32 + l] << 4));
    }
}

/* L/ggml/src/ggml-quants.c:1352-1374 */
void ora_dequantize_row_q4_K(const void *vx, float *y, int64_t k) {
    const ora_block_q4_K *x = (const ora_block_q4_K *)vx;
    const int64_t nb = k / ORA_QK_K;
    for (int64_t b = 0; b < nb; ++b) {
        uint8_t sc[8], mn[8];
        ora_unpack_scales_mins(x[b].scales, sc, mn);
        const float d = ora_fp16_to_fp32(x[b].d), dmin = ora_fp16_to_fp32(x[b].dmin);
        for (int g = 0; g < 4; ++g) {
            const float d_lo = d * sc[2 * g], m_lo = dmin * mn[2 * g];
            const float d_hi = d * sc[2 * g + 1], m_hi = dmin * mn[2 * g + 1];
            const uint8_t *q = x[b].qs + 32 * g;
            for (int l = 0; l < 32; ++l) y[64 * g + l]      = d_lo * (q[l] & 0x0f) - m_lo;
            for (int l = 0; l < 32; ++l) y[64 * g + 32 + l] = d_hi * (q[l] >> 4) - m_hi;
        }
        y += ORA_QK_K;
    }
}

/* ------------------------------------------------------------------------------------------------
 * the dot product — integers first (identical in every reference variant)
 * ---------------------------------------------------------------------------------------------- */

/* element e (0..255) of a superblock as the unsigned 4-bit level: low nibbles of the 32 bytes of
 * group g hold elements 64g..64g+31, high nibbles hold 64g+32..64g+63 (L/ggml/src/ggml-quants.c:1342-1346) */
static inline int q4_level(const ora_block_q4_K *x, int e) {
    const int g = e >> 6, r = e & 63;
    const uint8_t byte = x->qs[32 * g + (r & 31)];
    return r < 32 ? (byte & 0x0f) : (byte >> 4);
}

void ora_vec_dot_partials(int n, const void *vx, const void *vy, int32_t *p, int32_t *sumi, int32_t *summ) {
    const ora_block_q4_K *x = (const ora_block_q4_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    for (int b = 0; b < nb; ++b) {
        uint8_t sc[8], mn[8];
        ora_unpack_scales_mins(x[b].scales, sc, mn);
        int32_t si = 0, sm = 0;
        for (int j = 0; j < 8; ++j) {
            int32_t acc = 0;
            for (int l = 0; l < 32; ++l) acc += q4_level(&x[b], 32 * j + l) * (int32_t)y[b].qs[32 * j + l];
            if (p) p[8 * b + j] = acc;
            si += (int32_t)sc[j] * acc;
            sm += (int32_t)mn[j] * ((int32_t)y[b].bsums[2 * j] + (int32_t)y[b].bsums[2 * j + 1]);
        }
        if (sumi) sumi[b] = si;
        if (summ) summ[b] = sm;
    }
}

/* generic float order — L/ggml/src/ggml-cpu/quants.c:550-623.
 * aux32[l] (l = 0..7) collects scale * q4 * q8 over the elements with (e mod 8) == l; then
 * sums[l] += d * aux32[l]; sumf -= dmin * summ; result = sumf + sums[0] + ... + sums[7]. */
void ora_vec_dot_q4_K_q8_K_generic(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc) {
    (void)bs; (void)bx; (void)by; (void)nrc;
    const ora_block_q4_K *x = (const ora_block_q4_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    float lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float sumf = 0;
    for (int b = 0; b < nb; ++b) {
        uint8_t sc[8], mn[8];
        ora_unpack_scales_mins(x[b].scales, sc, mn);
        int32_t a32[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        int32_t sm = 0;
        for (int g = 0; g < 16; ++g) sm += (int32_t)y[b].bsums[g] * mn[g / 2];
        for (int e = 0; e < ORA_QK_K; ++e) {
            const int16_t prod = (int16_t)(y[b].qs[e] * q4_level(&x[b], e));
            a32[e & 7] += (int32_t)sc[e >> 5] * prod;
        }
        const float d = ora_fp16_to_fp32(x[b].d) * y[b].d;
        for (int l = 0; l < 8; ++l) lane[l] += d * (float)a32[l];
        const float dmin = ora_fp16_to_fp32(x[b].dmin) * y[b].d;
        sumf -= dmin * (float)sm;
    }
    for (int l = 0; l < 8; ++l) sumf += lane[l];
    *s = sumf;
}

/* AVX2 float order — L/ggml/src/ggml-cpu/arch/x86/quants.c:1760-1823, emulated lane by lane.
 *   32-bit lane t (0..7) of `sumi` collects, for every 64-element group g,
 *       sc[2g] * dot(elements 64g+4t .. +3)  +  sc[2g+1] * dot(elements 64g+32+4t .. +3)
 *   acc[t]   = fma(d,   (float)sumi[t], acc[t])               d    =  y.d * fp16(x.d)
 *   acc_m[u] = fma(dmin,(float)prod[u], acc_m[u])  u = 0..3   dmin = -y.d * fp16(x.dmin)
 *       prod[u] = m[2u]*(bsums[4u]+bsums[4u+1]) + m[2u+1]*(bsums[4u+2]+bsums[4u+3])
 *   result = hsum8(acc) + ((acc_m[0]+acc_m[2]) + (acc_m[1]+acc_m[3]))
 *   hsum8 (L/ggml/src/ggml-cpu/arch/x86/quants.c hsum_float_8): r[i]=a[i]+a[i+4]; r[0]+r[2], r[1]+r[3]; sum. */
void ora_vec_dot_q4_K_q8_K_avx2order(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc) {
    (void)bs; (void)bx; (void)by; (void)nrc;
    const ora_block_q4_K *x = (const ora_block_q4_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float acc_m[4] = {0, 0, 0, 0};
    for (int b = 0; b < nb; ++b) {
        const float d = y[b].d * ora_fp16_to_fp32(x[b].d);
        const float dmin = -y[b].d * ora_fp16_to_fp32(x[b].dmin);
        uint8_t sc[8], mn[8];
        ora_unpack_scales_mins(x[b].scales, sc, mn);
        for (int u = 0; u < 4; ++u) {
            const int32_t q0 = (int16_t)(y[b].bsums[4 * u] + y[b].bsums[4 * u + 1]);
            const int32_t q1 = (int16_t)(y[b].bsums[4 * u + 2] + y[b].bsums[4 * u + 3]);
            const int32_t prod = mn[2 * u] * q0 + mn[2 * u + 1] * q1;
            acc_m[u] = fmaf(dmin, (float)prod, acc_m[u]);
        }
        int32_t lane_sum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int g = 0; g < 4; ++g)
            for (int t = 0; t < 8; ++t) {
                int32_t lo = 0, hi = 0;
                for (int r = 0; r < 4; ++r) {
                    lo += q4_level(&x[b], 64 * g + 4 * t + r)      * (int32_t)y[b].qs[64 * g + 4 * t + r];
                    hi += q4_level(&x[b], 64 * g + 32 + 4 * t + r) * (int32_t)y[b].qs[64 * g + 32 + 4 * t + r];
                }
                lane_sum[t] += sc[2 * g] * lo + sc[2 * g + 1] * hi;
            }
        for (int t = 0; t < 8; ++t) acc[t] = fmaf(d, (float)lane_sum[t], acc[t]);
    }
    const float r0 = acc[0] + acc[4], r1 = acc[1] + acc[5], r2 = acc[2] + acc[6], r3 = acc[3] + acc[7];
    const float h = (r0 + r2) + (r1 + r3);
    const float hm = (acc_m[0] + acc_m[2]) + (acc_m[1] + acc_m[3]);
    *s = h + hm;
}

/* NEON float order — L/ggml/src/ggml-cpu/arch/arm/quants.c:2291-2351: one scalar accumulator,
 * per superblock `sumf -= dmin * summ` then `sumf += d * sumi`, unfused (SURVEY.md §3.4 notes that
 * aarch64 gcc may contract these; this restatement is the unfused C semantics). */
void ora_vec_dot_q4_K_q8_K_neonorder(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc) {
    (void)bs; (void)bx; (void)by; (void)nrc;
    const ora_block_q4_K *x = (const ora_block_q4_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    float sumf = 0;
    for (int b = 0; b < nb; ++b) {
        int32_t si, sm;
        ora_vec_dot_partials(ORA_QK_K, &x[b], &y[b], NULL, &si, &sm);
        const float d = y[b].d * ora_fp16_to_fp32(x[b].d);
        const float dmin = y[b].d * ora_fp16_to_fp32(x[b].dmin);
        sumf -= dmin * (float)sm;
        sumf += d * (float)si;
    }
    *s = sumf;
}

double ora_vec_dot_q4_K_q8_K_f64(int n, const void *vx, const void *vy) {
    const ora_block_q4_K *x = (const ora_block_q4_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    double acc = 0;
    for (int b = 0; b < nb; ++b) {
        int32_t si, sm;
        ora_vec_dot_partials(ORA_QK_K, &x[b], &y[b], NULL, &si, &sm);
        acc += (double)y[b].d * ((double)ora_fp16_to_fp32(x[b].d) * si - (double)ora_fp16_to_fp32(x[b].dmin) * sm);
    }
    return acc;
}

/* ------------------------------------------------------------------------------------------------
 * whole op — L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401: quantize every src1 row to Q8_K into wdata
 * (:1267-1307), then dst[i1*nb1 + ir0*4] = vec_dot(ne00, src0 row ir0, wdata row i1) (:1192-1195).
 * The reference's threading / chunking does not change any value, so it is not restated.
 * ---------------------------------------------------------------------------------------------- */
int ora_mul_mat_q4_K_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride,
                         const float *x, int64_t N, int64_t x_row_stride_bytes,
                         float *dst, int64_t dst_col_stride_bytes, void *wdata, int order) {
    if (K <= 0 || K % ORA_QK_K != 0 || M < 0 || N < 0) return -1;
    const int64_t q8_row = K / ORA_QK_K * (int64_t)sizeof(ora_block_q8_K);
    for (int64_t c = 0; c < N; ++c) {
        /* the reference leaves bsums of all-zero blocks unwritten; zero them here so the result is defined */
        memset((char *)wdata + c * q8_row, 0, (size_t)q8_row);
        ora_quantize_row_q8_K((const float *)((const char *)x + c * x_row_stride_bytes), (char *)wdata + c * q8_row, K);
    }
    for (int64_t c = 0; c < N; ++c) {
        float *out = (float *)((char *)dst + c * dst_col_stride_bytes);
        const void *yq = (const char *)wdata + c * q8_row;
        for (int64_t r = 0; r < M; ++r) {
            const void *xr = (const char *)w + r * w_row_stride;
            switch (order) {
                case ORA_ORDER_AVX2: ora_vec_dot_q4_K_q8_K_avx2order((int)K, out + r, 0, xr, 0, yq, 0, 1); break;
                case ORA_ORDER_NEON: ora_vec_dot_q4_K_q8_K_neonorder((int)K, out + r, 0, xr, 0, yq, 0, 1); break;
                default:             ora_vec_dot_q4_K_q8_K_generic((int)K, out + r, 0, xr, 0, yq, 0, 1); break;
            }
        }
    }
    return 0;
}
