/*
 * q6k_oracle.c — CPU ORACLE for the Q6_K x Q8_K mul_mat path (test infrastructure, see q4k_oracle.h).
 *
 * TinyLlama Q4_K_M keeps attn_v / ffn_down of 10 layers and output.weight in Q6_K
 * (L/src/llama-quant.cpp:228-230, 302-303, 358-364), so the decode token needs this dot product too.
 * Plain-C restatement (not a copy); every function cites the reference lines it follows.  Built with
 * -ffp-contract=off like q4k_oracle.c; the one fused operation of the AVX2 branch is an explicit fmaf().
 *
 * "L/" = /root/reference/external/llama.cpp/
 */
#include "q4k_oracle.h"

#include <math.h>
#include <string.h>

_Static_assert(sizeof(ora_block_q6_K) == 210, "block_q6_K is 210 bytes (L/ggml/src/ggml-common.h:336)");
_Static_assert(offsetof(ora_block_q6_K, qh) == 128 && offsetof(ora_block_q6_K, scales) == 192 &&
               offsetof(ora_block_q6_K, d) == 208, "q6_K offsets");

/* the signed 6-bit level of element e (0..255) of a block: L/ggml/src/ggml-quants.c:1772-1779 (dequantize_row_q6_K)
 * and L/ggml/src/ggml-cpu/quants.c:731-738.  Element 128h + 32t + l lives in nibble (t >> 1) of ql[64h + 32(t&1) + l]
 * and in bits 2t..2t+1 of qh[32h + l]. */
static inline int q6_level(const ora_block_q6_K *x, int e) {
    const int h = e >> 7, t = (e >> 5) & 3, l = e & 31;
    const uint8_t lo = x->ql[64 * h + 32 * (t & 1) + l];
    const int nib = (t & 2) ? (lo >> 4) : (lo & 0xF);
    const int top = (x->qh[32 * h + l] >> (2 * t)) & 3;
    return (nib | (top << 4)) - 32;
}

/* ------------------------------------------------------------------------------------------------
 * weight quantizer — L/ggml/src/ggml-quants.c:451-517 (make_qx_quants, rmse_type 1, no weights) and
 * :1692-1760 (quantize_row_q6_K_ref).  Input generator only (never accelerated).
 * ---------------------------------------------------------------------------------------------- */
static int clamp6(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

static float search_scale_q6(const float *x, int8_t *L) {     /* n = 16, nmax = 32 */
    const int n = 16, nmax = 32;
    float max = 0, amax = 0;
    for (int i = 0; i < n; ++i) {
        const float ax = fabsf(x[i]);
        if (ax > amax) { amax = ax; max = x[i]; }
    }
    if (amax < 1e-15f) {                                        /* GROUP_MAX_EPS, :16 */
        for (int i = 0; i < n; ++i) L[i] = 0;
        return 0.0f;
    }
    float iscale = -nmax / max;
    float sumlx = 0, suml2 = 0;
    for (int i = 0; i < n; ++i) {
        const int l = clamp6(ora_nearest_int(iscale * x[i]), -nmax, nmax - 1);
        L[i] = (int8_t)(l + nmax);
        const float w = x[i] * x[i];
        sumlx += w * x[i] * l;
        suml2 += w * l * l;
    }
    float scale = suml2 ? sumlx / suml2 : 0.0f;
    float best = scale * sumlx;
    for (int is = -9; is <= 9; ++is) {
        if (is == 0) continue;
        iscale = -(nmax + 0.1f * is) / max;
        sumlx = suml2 = 0;
        for (int i = 0; i < n; ++i) {
            const int l = clamp6(ora_nearest_int(iscale * x[i]), -nmax, nmax - 1);
            const float w = x[i] * x[i];
            sumlx += w * x[i] * l;
            suml2 += w * l * l;
        }
        if (suml2 > 0 && sumlx * sumlx > best * suml2) {
            for (int i = 0; i < n; ++i) L[i] = (int8_t)(nmax + clamp6(ora_nearest_int(iscale * x[i]), -nmax, nmax - 1));
            scale = sumlx / suml2;
            best = scale * sumlx;
        }
    }
    return scale;
}

void ora_quantize_row_q6_K(const float *x, void *vy, int64_t k) {
    ora_block_q6_K *y = (ora_block_q6_K *)vy;
    const int64_t nb = k / ORA_QK_K;
    int8_t L[ORA_QK_K];
    float scales[16];
    for (int64_t b = 0; b < nb; ++b, x += ORA_QK_K) {
        float max_scale = 0, max_abs_scale = 0;
        for (int ib = 0; ib < 16; ++ib) {
            const float scale = search_scale_q6(x + 16 * ib, L + 16 * ib);
            scales[ib] = scale;
            const float a = fabsf(scale);
            if (a > max_abs_scale) { max_abs_scale = a; max_scale = scale; }
        }
        if (max_abs_scale < 1e-15f) {                           /* :1715-1720 */
            memset(&y[b], 0, sizeof y[b]);
            continue;
        }
        const float iscale = -128.0f / max_scale;
        y[b].d = ora_fp32_to_fp16(1 / iscale);
        const float dd = ora_fp16_to_fp32(y[b].d);
        for (int ib = 0; ib < 16; ++ib) {
            const int v = ora_nearest_int(iscale * scales[ib]);
            y[b].scales[ib] = (int8_t)(v < 127 ? v : 127);
        }
        for (int j = 0; j < 16; ++j) {
            const float d = dd * y[b].scales[j];
            if (!d) continue;                                   /* L keeps the search's levels, :1731-1733 */
            for (int ii = 0; ii < 16; ++ii)
                L[16 * j + ii] = (int8_t)(clamp6(ora_nearest_int(x[16 * j + ii] / d), -32, 31) + 32);
        }
        for (int h = 0; h < 2; ++h)
            for (int l = 0; l < 32; ++l) {
                const int8_t *Lh = L + 128 * h + l;
                y[b].ql[64 * h + l]      = (uint8_t)((Lh[0] & 0xF) | ((Lh[64] & 0xF) << 4));
                y[b].ql[64 * h + 32 + l] = (uint8_t)((Lh[32] & 0xF) | ((Lh[96] & 0xF) << 4));
                y[b].qh[32 * h + l] = (uint8_t)((Lh[0] >> 4) | ((Lh[32] >> 4) << 2) | ((Lh[64] >> 4) << 4) | ((Lh[96] >> 4) << 6));
            }
    }
}

/* L/ggml/src/ggml-quants.c:1762-1791 — y = d * scales[e / 16] * level(e), evaluated as (d * sc) * q */
void ora_dequantize_row_q6_K(const void *vx, float *y, int64_t k) {
    const ora_block_q6_K *x = (const ora_block_q6_K *)vx;
    const int64_t nb = k / ORA_QK_K;
    for (int64_t b = 0; b < nb; ++b, y += ORA_QK_K) {
        const float d = ora_fp16_to_fp32(x[b].d);
        for (int e = 0; e < ORA_QK_K; ++e) y[e] = d * x[b].scales[e >> 4] * (float)q6_level(&x[b], e);
    }
}

/* ------------------------------------------------------------------------------------------------
 * integer stage: p[nb*16] = dot of the 16 levels of sub-block j with the Q8_K bytes (levels already offset by -32),
 * sumi[nb] = sum_j scales[j] * p[j].  Every variant of the reference computes these integers
 * (generic: aux32, L/ggml/src/ggml-cpu/quants.c:740-750; AVX2: maddubs/sub/madd, arch/x86/quants.c:2190-2210).
 * ---------------------------------------------------------------------------------------------- */
void ora_vec_dot_q6_partials(int n, const void *vx, const void *vy, int32_t *p, int32_t *sumi) {
    const ora_block_q6_K *x = (const ora_block_q6_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    for (int b = 0; b < nb; ++b) {
        int32_t total = 0;
        for (int j = 0; j < 16; ++j) {
            int32_t s = 0;
            for (int t = 0; t < 16; ++t) s += q6_level(&x[b], 16 * j + t) * (int32_t)y[b].qs[16 * j + t];
            if (p) p[b * 16 + j] = s;
            total += x[b].scales[j] * s;
        }
        if (sumi) sumi[b] = total;
    }
}

/* generic float order — L/ggml/src/ggml-cpu/quants.c:705-755: eight running sums indexed by element mod 8,
 * `sums[l] += d * aux32[l]` per superblock (unfused), then summed 0..7. */
void ora_vec_dot_q6_K_q8_K_generic(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc) {
    (void)bs; (void)bx; (void)by; (void)nrc;
    const ora_block_q6_K *x = (const ora_block_q6_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    float sums[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int b = 0; b < nb; ++b) {
        int32_t aux32[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int e = 0; e < ORA_QK_K; ++e)
            aux32[e & 7] += x[b].scales[e >> 4] * (int32_t)(int16_t)(q6_level(&x[b], e) * y[b].qs[e]);
        const float d = ora_fp16_to_fp32(x[b].d) * y[b].d;
        for (int l = 0; l < 8; ++l) sums[l] += d * (float)aux32[l];
    }
    float sumf = 0;
    for (int l = 0; l < 8; ++l) sumf += sums[l];
    *s = sumf;
}

/* AVX2 float order — L/ggml/src/ggml-cpu/arch/x86/quants.c:2145-2218: int32 lane m (0..7) of `sumi` collects, for
 * every 32-element group g, scale(2g + (m >> 2)) * sum of elements 4m..4m+3 of the group (maddubs pairs, minus
 * 32 * q8 pairs, madd_epi16 with the sign-extended scales); acc = fmadd(d, (float)sumi, acc) per superblock with
 * d = y.d * fp16(x.d); hsum_float_8 at the end. */
void ora_vec_dot_q6_K_q8_K_avx2order(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc) {
    (void)bs; (void)bx; (void)by; (void)nrc;
    const ora_block_q6_K *x = (const ora_block_q6_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int b = 0; b < nb; ++b) {
        const float d = y[b].d * ora_fp16_to_fp32(x[b].d);
        int32_t lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int g = 0; g < 8; ++g)
            for (int m = 0; m < 8; ++m) {
                int32_t t = 0;
                for (int r = 0; r < 4; ++r) t += q6_level(&x[b], 32 * g + 4 * m + r) * (int32_t)y[b].qs[32 * g + 4 * m + r];
                lane[m] += x[b].scales[2 * g + (m >> 2)] * t;
            }
        for (int m = 0; m < 8; ++m) acc[m] = fmaf(d, (float)lane[m], acc[m]);
    }
    const float r0 = acc[0] + acc[4], r1 = acc[1] + acc[5], r2 = acc[2] + acc[6], r3 = acc[3] + acc[7];
    *s = (r0 + r2) + (r1 + r3);
}

/* same integers, combined in double: the tie-breaker for tolerance tests */
double ora_vec_dot_q6_K_q8_K_f64(int n, const void *vx, const void *vy) {
    const ora_block_q6_K *x = (const ora_block_q6_K *)vx;
    const ora_block_q8_K *y = (const ora_block_q8_K *)vy;
    const int nb = n / ORA_QK_K;
    double acc = 0;
    for (int b = 0; b < nb; ++b) {
        int32_t si;
        ora_vec_dot_q6_partials(ORA_QK_K, &x[b], &y[b], NULL, &si);
        acc += (double)y[b].d * (double)ora_fp16_to_fp32(x[b].d) * si;
    }
    return acc;
}

/* whole op, as ora_mul_mat_q4_K_f32 (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401 with type_traits_cpu[GGML_TYPE_Q6_K]:
 * vec_dot = ggml_vec_dot_q6_K_q8_K, vec_dot_type = Q8_K, :296-301) */
int ora_mul_mat_q6_K_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride,
                         const float *x, int64_t N, int64_t x_row_stride_bytes,
                         float *dst, int64_t dst_col_stride_bytes, void *wdata, int order) {
    if (K <= 0 || K % ORA_QK_K != 0 || M < 0 || N < 0) return -1;
    const int64_t q8_row = K / ORA_QK_K * (int64_t)sizeof(ora_block_q8_K);
    for (int64_t c = 0; c < N; ++c) {
        memset((char *)wdata + c * q8_row, 0, (size_t)q8_row);
        ora_quantize_row_q8_K((const float *)((const char *)x + c * x_row_stride_bytes), (char *)wdata + c * q8_row, K);
    }
    for (int64_t c = 0; c < N; ++c) {
        float *out = (float *)((char *)dst + c * dst_col_stride_bytes);
        const void *yq = (const char *)wdata + c * q8_row;
        for (int64_t r = 0; r < M; ++r) {
            const void *xr = (const char *)w + r * w_row_stride;
            if (order == ORA_ORDER_AVX2) ora_vec_dot_q6_K_q8_K_avx2order((int)K, out + r, 0, xr, 0, yq, 0, 1);
            else                         ora_vec_dot_q6_K_q8_K_generic((int)K, out + r, 0, xr, 0, yq, 0, 1);
        }
    }
    return 0;
}
