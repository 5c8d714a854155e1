/*
 * q4k_oracle.h — CPU ORACLE for the Q4_K x Q8_K mul_mat hot path.
 *
 * THIS IS TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product path (ggml-neon-opt_b200/) never links or calls it.
 *
 * It is a plain-C restatement (not a copy) of the reference's algorithm; every function cites the
 * reference file:line it follows ("L/" = /root/reference/external/llama.cpp/).  It is pinned against
 * the unmodified reference compiled into oracle/_ref/ (tests/test_oracle_vs_ref.py, run in the build
 * container) and against fixtures generated from that reference (tests/golden/, tests/test_golden.py).
 * The reference holds no golden vectors of its own for this path.
 */
#ifndef Q4K_ORACLE_H
#define Q4K_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORA_QK_K 256          /* L/ggml/src/ggml-common.h:89  */
#define ORA_K_SCALE_SIZE 12   /* L/ggml/src/ggml-common.h:90  */

/* L/ggml/src/ggml-common.h:291-306 — 144 bytes */
typedef struct {
    uint16_t d;                         /* fp16 super-block scale for the 6-bit scales */
    uint16_t dmin;                      /* fp16 super-block scale for the 6-bit mins   */
    uint8_t  scales[ORA_K_SCALE_SIZE];  /* 8 x 6-bit scale + 8 x 6-bit min             */
    uint8_t  qs[ORA_QK_K / 2];          /* 4-bit quants                                */
} ora_block_q4_K;

/* L/ggml/src/ggml-common.h:338-344 — 292 bytes */
typedef struct {
    float   d;
    int8_t  qs[ORA_QK_K];
    int16_t bsums[ORA_QK_K / 16];
} ora_block_q8_K;

/* L/ggml/src/ggml-common.h:326-336 — 210 bytes */
typedef struct {
    uint8_t  ql[ORA_QK_K / 2];          /* low 4 bits of the 6-bit levels  */
    uint8_t  qh[ORA_QK_K / 4];          /* high 2 bits                     */
    int8_t   scales[ORA_QK_K / 16];     /* 16 signed 8-bit sub-block scales */
    uint16_t d;                         /* fp16 super-block scale          */
} ora_block_q6_K;

/* float accumulation orders of the reference's variants */
enum ora_float_order {
    ORA_ORDER_GENERIC = 0,   /* L/ggml/src/ggml-cpu/quants.c:550-623            */
    ORA_ORDER_AVX2    = 1,   /* L/ggml/src/ggml-cpu/arch/x86/quants.c:1760-1823 */
    ORA_ORDER_NEON    = 2,   /* L/ggml/src/ggml-cpu/arch/arm/quants.c:2291-2351 */
};

int      ora_nearest_int(float f);
uint16_t ora_fp32_to_fp16(float f);
float    ora_fp16_to_fp32(uint16_t h);

void ora_unpack_scales_mins(const uint8_t packed[12], uint8_t sc[8], uint8_t mn[8]);

void ora_quantize_row_q8_K(const float *x, void *y, int64_t k);
void ora_quantize_row_q4_K(const float *x, void *y, int64_t k);
void ora_dequantize_row_q4_K(const void *x, float *y, int64_t k);

/* integer stage: p[nb*8] per-32 sub-block dot products, sumi[nb], summ[nb] */
void ora_vec_dot_partials(int n, const void *vx, const void *vy, int32_t *p, int32_t *sumi, int32_t *summ);

/* ggml_vec_dot_t signature (L/ggml/include/ggml-cpu.h:109-117); nrc must be 1 */
void ora_vec_dot_q4_K_q8_K_generic(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
void ora_vec_dot_q4_K_q8_K_avx2order(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
void ora_vec_dot_q4_K_q8_K_neonorder(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
/* same integers, final combination in double: the tie-breaker "true" value for tolerance tests */
double ora_vec_dot_q4_K_q8_K_f64(int n, const void *vx, const void *vy);

/* whole op: dst[n*M + m] = vec_dot(K, W row m, quantize_q8_K(X row n)); L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401 */
int ora_mul_mat_q4_K_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride,
                         const float *x, int64_t N, int64_t x_row_stride_bytes,
                         float *dst, int64_t dst_col_stride_bytes,
                         void *wdata /* N * K/256 * 292 bytes */, int order);

/* ---- Q6_K x Q8_K (q6k_oracle.c) ---- */
void ora_quantize_row_q6_K(const float *x, void *y, int64_t k);       /* L/ggml/src/ggml-quants.c:1692-1760 */
void ora_dequantize_row_q6_K(const void *x, float *y, int64_t k);     /* L/ggml/src/ggml-quants.c:1762-1791 */
/* integer stage: p[nb*16] per-16 sub-block dot products (levels offset by -32), sumi[nb] = sum_j scales[j]*p[j] */
void ora_vec_dot_q6_partials(int n, const void *vx, const void *vy, int32_t *p, int32_t *sumi);
void ora_vec_dot_q6_K_q8_K_generic(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
void ora_vec_dot_q6_K_q8_K_avx2order(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
double ora_vec_dot_q6_K_q8_K_f64(int n, const void *vx, const void *vy);
int ora_mul_mat_q6_K_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride,
                         const float *x, int64_t N, int64_t x_row_stride_bytes,
                         float *dst, int64_t dst_col_stride_bytes, void *wdata, int order);

#ifdef __cplusplus
}
#endif
#endif
