/*
 * ref_shim.c — TEST INFRASTRUCTURE.  A few entry points that drive the UNMODIFIED reference
 * (oracle/_ref/libggml*.so, built by oracle/Makefile from /root/reference) through its own public
 * API, so Python (ctypes) can run the reference's whole op and load our backend plugin the same way
 * llama.cpp tools would.  Compiled against the reference's headers where they lie; links only
 * against the reference's libraries.  Nothing from the product path is linked here.
 *
 *   ref_mul_mat_q4k        ggml_mul_mat(Q4_K[K,M], F32[K,N]) on ggml_backend_cpu (ggml-cpu.c:1205)
 *   ref_plugin_mul_mat_q4k same graph on a backend loaded with ggml_backend_load(path)
 *                          (ggml-backend-reg.cpp:247-291, the GGML_BACKEND_PATH mechanism)
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

typedef struct {
    struct ggml_context *ctx;
    struct ggml_tensor *w, *x, *y;
    struct ggml_cgraph *gf;
    ggml_backend_buffer_t buf;
} shim_graph;

static int build_graph(shim_graph *g, ggml_backend_t backend, ggml_backend_buffer_type_t buft,
                       int64_t M, int64_t K, int64_t N) {
    struct ggml_init_params ip = { ggml_tensor_overhead() * 8 + ggml_graph_overhead(), NULL, true };
    g->ctx = ggml_init(ip);
    if (!g->ctx) return -1;
    g->w = ggml_new_tensor_2d(g->ctx, GGML_TYPE_Q4_K, K, M);
    g->x = ggml_new_tensor_2d(g->ctx, GGML_TYPE_F32, K, N);
    g->y = ggml_mul_mat(g->ctx, g->w, g->x);
    g->gf = ggml_new_graph(g->ctx);
    ggml_build_forward_expand(g->gf, g->y);
    g->buf = buft ? ggml_backend_alloc_ctx_tensors_from_buft(g->ctx, buft)
                  : ggml_backend_alloc_ctx_tensors(g->ctx, backend);
    return g->buf ? 0 : -2;
}

static void free_graph(shim_graph *g) {
    if (g->buf) ggml_backend_buffer_free(g->buf);
    if (g->ctx) ggml_free(g->ctx);
}

/* Times `iters` graph computes after `warmup`; returns 0 on success.  secs[0]=best, secs[1]=mean. */
int ref_mul_mat_q4k(const void *w_q4k, int64_t M, int64_t K, const float *x, int64_t N, float *dst,
                    int n_threads, int warmup, int iters, double *secs) {
    ggml_backend_t cpu = ggml_backend_cpu_init();
    if (!cpu) return -1;
    ggml_backend_cpu_set_n_threads(cpu, n_threads);
    shim_graph g = {0};
    /* plain CPU buffer: keeps the weights in block_q4_K layout so the vec_dot path is taken,
     * not the CPU_REPACK extra buffer type (L/ggml/src/ggml-cpu/repack.cpp:1844-1849) */
    int rc = build_graph(&g, cpu, ggml_backend_cpu_buffer_type(), M, K, N);
    if (rc) { free_graph(&g); ggml_backend_free(cpu); return rc; }
    ggml_backend_tensor_set(g.w, w_q4k, 0, ggml_nbytes(g.w));
    ggml_backend_tensor_set(g.x, x, 0, ggml_nbytes(g.x));
    for (int i = 0; i < warmup; ++i) ggml_backend_graph_compute(cpu, g.gf);
    double best = 1e30, total = 0;
    for (int i = 0; i < iters; ++i) {
        const double t0 = now_s();
        if (ggml_backend_graph_compute(cpu, g.gf) != GGML_STATUS_SUCCESS) { rc = -3; break; }
        const double dt = now_s() - t0;
        total += dt;
        if (dt < best) best = dt;
    }
    if (secs) { secs[0] = best; secs[1] = iters > 0 ? total / iters : 0; }
    if (dst) ggml_backend_tensor_get(g.y, dst, 0, ggml_nbytes(g.y));
    free_graph(&g);
    ggml_backend_free(cpu);
    return rc;
}

/* Loads the backend plugin at `path` through the reference's registry and runs the same 1-op graph
 * on its first device, with all three tensors in the device's default buffer type (what
 * test-backend-ops does, L/tests/test-backend-ops.cpp:1180).  Returns 0 on success, <0 on error:
 * -10 load failed, -11 no device, -12 supports_op says no, -13 init failed. */
int ref_plugin_mul_mat_q4k(const char *path, const void *w_q4k, int64_t M, int64_t K, const float *x,
                           int64_t N, float *dst, int iters, double *secs) {
    static ggml_backend_reg_t reg = NULL;   /* the registry never unloads a backend (ggml-backend-reg.cpp:215-223) */
    if (!reg) reg = ggml_backend_load(path);
    if (!reg) return -10;
    if (ggml_backend_reg_dev_count(reg) < 1) return -11;
    ggml_backend_dev_t dev = ggml_backend_reg_dev_get(reg, 0);
    ggml_backend_t be = ggml_backend_dev_init(dev, NULL);
    if (!be) return -13;
    shim_graph g = {0};
    int rc = build_graph(&g, be, NULL, M, K, N);
    if (rc) { free_graph(&g); ggml_backend_free(be); return rc; }
    if (!ggml_backend_dev_supports_op(dev, g.y)) { free_graph(&g); ggml_backend_free(be); return -12; }
    ggml_backend_tensor_set(g.w, w_q4k, 0, ggml_nbytes(g.w));
    double best = 1e30, total = 0;
    for (int i = 0; i < (iters > 0 ? iters : 1); ++i) {
        const double t0 = now_s();
        ggml_backend_tensor_set(g.x, x, 0, ggml_nbytes(g.x));          /* split-input copy, ggml-backend.cpp:1427-1450 */
        if (ggml_backend_graph_compute(be, g.gf) != GGML_STATUS_SUCCESS) { rc = -3; break; }
        ggml_backend_tensor_get(g.y, dst, 0, ggml_nbytes(g.y));
        const double dt = now_s() - t0;
        total += dt;
        if (dt < best) best = dt;
    }
    if (secs) { secs[0] = best; secs[1] = total / (iters > 0 ? iters : 1); }
    free_graph(&g);
    ggml_backend_free(be);
    return rc;
}

/* n_mats Q4_K mul_mats on ONE activation tensor (attn_q / attn_k / attn_v of a layer: three ggml_mul_mat nodes on the same
 * `cur`), as one graph on the plugin's first device.  w_q4k[i] is [Ms[i]][K/256*144], dst[i] receives [N][Ms[i]]. */
int ref_plugin_mul_mat_shared_q4k(const char *path, int n_mats, const void *const *w_q4k, const int64_t *Ms, int64_t K,
                                  const float *x, int64_t N, float *const *dst) {
    static ggml_backend_reg_t reg = NULL;
    if (!reg) reg = ggml_backend_load(path);
    if (!reg) return -10;
    if (ggml_backend_reg_dev_count(reg) < 1) return -11;
    ggml_backend_dev_t dev = ggml_backend_reg_dev_get(reg, 0);
    ggml_backend_t be = ggml_backend_dev_init(dev, NULL);
    if (!be) return -13;
    struct ggml_init_params ip = { ggml_tensor_overhead() * (size_t)(2 * n_mats + 8) + ggml_graph_overhead(), NULL, true };
    struct ggml_context *ctx = ggml_init(ip);
    if (!ctx) { ggml_backend_free(be); return -1; }
    struct ggml_cgraph *gf = ggml_new_graph(ctx);
    struct ggml_tensor *xs = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, N);
    struct ggml_tensor *ws[8], *ys[8];
    int rc = 0;
    if (n_mats > 8) { rc = -1; goto out; }
    for (int i = 0; i < n_mats; ++i) {
        ws[i] = ggml_new_tensor_2d(ctx, GGML_TYPE_Q4_K, K, Ms[i]);
        ys[i] = ggml_mul_mat(ctx, ws[i], xs);
        ggml_build_forward_expand(gf, ys[i]);
    }
    {
        ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, be);
        if (!buf) { rc = -2; goto out; }
        for (int i = 0; i < n_mats; ++i) {
            if (!ggml_backend_dev_supports_op(dev, ys[i])) { rc = -12; break; }
            ggml_backend_tensor_set(ws[i], w_q4k[i], 0, ggml_nbytes(ws[i]));
        }
        if (!rc) {
            ggml_backend_tensor_set(xs, x, 0, ggml_nbytes(xs));
            if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) rc = -3;
            for (int i = 0; i < n_mats && !rc; ++i) ggml_backend_tensor_get(ys[i], dst[i], 0, ggml_nbytes(ys[i]));
        }
        ggml_backend_buffer_free(buf);
    }
out:
    ggml_free(ctx);
    ggml_backend_free(be);
    return rc;
}

int ref_cpu_has_avx2(void) { return ggml_cpu_has_avx2(); }
int ref_n_backend_devices(void) { return (int)ggml_backend_dev_count(); }

/* A whole set of independent Q4_K mul_mats in ONE graph (bench.py's reference arm: the Q4_K matmuls of one
 * decode token).  Weights are filled with valid random block_q4_K bytes (xorshift; fp16 d = 2^-4,
 * dmin = 2^-5 — throughput only, the parity tests use real quantized weights), activations with a
 * deterministic ramp.  Times `iters` graph computes after `warmup`.  secs[0] = best, secs[1] = mean.
 * checksum (optional) receives the sum of all outputs of the last compute. */
/* same with a weight type per matrix: q6[i] != 0 -> GGML_TYPE_Q6_K (random valid block_q6_K bytes, d = 2^-8), else Q4_K:
 * the full Q4_K_M decode token */
/* random valid block bytes (xorshift; Q4_K: fp16 d = 2^-4, dmin = 2^-5; Q6_K: d = 2^-8) */
static void fill_random_blocks(uint8_t *p, size_t nbytes, int q6, uint64_t *state) {
    uint64_t s = *state;
    for (size_t o = 0; o + 8 <= nbytes; o += 8) {
        s ^= s << 13; s ^= s >> 7; s ^= s << 17;
        memcpy(p + o, &s, 8);
    }
    if (q6) { for (size_t o = 0; o < nbytes; o += 210) { p[o + 208] = 0x00; p[o + 209] = 0x1c; } }
    else { for (size_t o = 0; o < nbytes; o += 144) { p[o] = 0x00; p[o + 1] = 0x2c; p[o + 2] = 0x00; p[o + 3] = 0x28; } }
    *state = s;
}

/* the CPU backend's CPU_REPACK extra buffer type (what llama's loader puts Q4_K weights in on an AVX2 host:
 * L/ggml/src/ggml-cpu/repack.cpp:1844-1849 -> ggml_gemv/gemm_q4_K_8x8_q8_K), reached the way llama does: through the
 * registry's "ggml_backend_dev_get_extra_bufts" proc address (L/ggml/src/ggml-cpu/ggml-cpu.cpp:624-632) */
static ggml_backend_buffer_type_t cpu_repack_buft(ggml_backend_t cpu) {
    ggml_backend_dev_t dev = ggml_backend_get_device(cpu);
    ggml_backend_reg_t reg = ggml_backend_dev_backend_reg(dev);
    typedef ggml_backend_buffer_type_t *(*get_extra_t)(ggml_backend_dev_t);
    get_extra_t fn = (get_extra_t)ggml_backend_reg_get_proc_address(reg, "ggml_backend_dev_get_extra_bufts");
    if (!fn) return NULL;
    for (ggml_backend_buffer_type_t *b = fn(dev); b && *b; ++b)
        if (strstr(ggml_backend_buft_name(*b), "REPACK")) return *b;
    return NULL;
}

/* repack != 0: the weights live in the CPU_REPACK buffer type (repacked to block_q4_Kx8 by its set_tensor), as llama-bench runs
 * them by default on an AVX2 host; Q6_K has no repack format on AVX2 and stays plain.  -20: no repack buffer type in this build. */
int ref_mul_mat_set_ex(int n_mats, const int64_t *Ms, const int64_t *Ks, const int *q6, int64_t N, int n_threads,
                       int warmup, int iters, int repack, double *secs, double *checksum) {
    ggml_backend_t cpu = ggml_backend_cpu_init();
    if (!cpu) return -1;
    ggml_backend_cpu_set_n_threads(cpu, n_threads);
    struct ggml_init_params ip = { ggml_tensor_overhead() * (size_t)(3 * n_mats + 8) +
                                   ggml_graph_overhead_custom((size_t)(4 * n_mats + 16), false), NULL, true };
    struct ggml_init_params ipw = { ggml_tensor_overhead() * (size_t)(n_mats + 8), NULL, true };
    struct ggml_context *ctx = ggml_init(ip), *wctx = ggml_init(ipw);
    if (!ctx || !wctx) { ggml_backend_free(cpu); return -1; }
    struct ggml_cgraph *gf = ggml_new_graph_custom(ctx, (size_t)(4 * n_mats + 16), false);
    struct ggml_tensor **ws = calloc((size_t)n_mats, sizeof *ws), **xs = calloc((size_t)n_mats, sizeof *xs),
                       **ys = calloc((size_t)n_mats, sizeof *ys);
    for (int i = 0; i < n_mats; ++i) {
        /* repacked weights sit in their own context / buffer; everything else in a plain CPU buffer */
        ws[i] = ggml_new_tensor_2d(repack && !(q6 && q6[i]) ? wctx : ctx, q6 && q6[i] ? GGML_TYPE_Q6_K : GGML_TYPE_Q4_K, Ks[i], Ms[i]);
        xs[i] = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, Ks[i], N);
        ys[i] = ggml_mul_mat(ctx, ws[i], xs[i]);
        ggml_build_forward_expand(gf, ys[i]);
    }
    int rc = 0;
    ggml_backend_buffer_t wbuf = NULL;
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors_from_buft(ctx, ggml_backend_cpu_buffer_type());
    if (!buf) { rc = -2; goto done; }
    if (repack) {
        ggml_backend_buffer_type_t rb = cpu_repack_buft(cpu);
        if (!rb) { rc = -20; goto done; }
        wbuf = ggml_backend_alloc_ctx_tensors_from_buft(wctx, rb);
        if (!wbuf) { rc = -2; goto done; }
    }
    uint64_t s = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < n_mats; ++i) {
        const size_t nbytes = ggml_nbytes(ws[i]);
        const int is6 = q6 && q6[i];
        if (repack && !is6) {                      /* through set_tensor: that is where the repack happens */
            uint8_t *tmp = malloc(nbytes);
            if (!tmp) { rc = -1; goto done; }
            fill_random_blocks(tmp, nbytes, 0, &s);
            ggml_backend_tensor_set(ws[i], tmp, 0, nbytes);
            free(tmp);
        } else {
            fill_random_blocks((uint8_t *)ws[i]->data, nbytes, is6, &s);
        }
        float *x = (float *)xs[i]->data;
        const int64_t ne = Ks[i] * N;
        for (int64_t j = 0; j < ne; ++j) x[j] = (float)((j * 2654435761u) % 2001) * 1e-3f - 1.0f;
    }
    for (int i = 0; i < warmup; ++i) ggml_backend_graph_compute(cpu, gf);
    double best = 1e30, total = 0;
    for (int i = 0; i < iters; ++i) {
        const double t0 = now_s();
        if (ggml_backend_graph_compute(cpu, gf) != GGML_STATUS_SUCCESS) { rc = -3; break; }
        const double dt = now_s() - t0;
        total += dt;
        if (dt < best) best = dt;
    }
    if (secs) { secs[0] = best; secs[1] = iters > 0 ? total / iters : 0; }
    if (checksum) {
        double c = 0;
        for (int i = 0; i < n_mats; ++i) {
            const float *y = (const float *)ys[i]->data;
            for (int64_t j = 0; j < Ms[i] * N; ++j) c += y[j];
        }
        *checksum = c;
    }
done:
    if (buf) ggml_backend_buffer_free(buf);
    if (wbuf) ggml_backend_buffer_free(wbuf);
    ggml_free(ctx);
    ggml_free(wctx);
    free(ws); free(xs); free(ys);
    ggml_backend_free(cpu);
    return rc;
}

int ref_mul_mat_set_typed(int n_mats, const int64_t *Ms, const int64_t *Ks, const int *q6, int64_t N, int n_threads,
                          int warmup, int iters, double *secs, double *checksum) {
    return ref_mul_mat_set_ex(n_mats, Ms, Ks, q6, N, n_threads, warmup, iters, 0, secs, checksum);
}

/* BASELINE.md §5 timing (i), "direct": quantize_row_q8_K of the N activation rows + one ggml_vec_dot_q4_K_q8_K per (row, column)
 * on ONE thread, through the reference's own type-traits table (L/ggml/src/ggml-cpu/ggml-cpu.c:274-283, 355-357) — the loop
 * ggml_compute_forward_mul_mat_one_chunk runs, without the threadpool.  Only rows [0, m_sample) are computed (bounded CPU
 * time); weights random valid blocks.  secs[0] = best, secs[1] = mean seconds per pass over the sample. */
int ref_direct_q4k(int64_t m_sample, int64_t K, int64_t N, int warmup, int iters, double *secs, double *checksum) {
    const struct ggml_type_traits_cpu *tq4 = ggml_get_type_traits_cpu(GGML_TYPE_Q4_K), *tq8 = ggml_get_type_traits_cpu(GGML_TYPE_Q8_K);
    if (!tq4 || !tq4->vec_dot || !tq8 || !tq8->from_float || tq4->vec_dot_type != GGML_TYPE_Q8_K) return -1;
    const size_t wrow = ggml_row_size(GGML_TYPE_Q4_K, K), yrow = ggml_row_size(GGML_TYPE_Q8_K, K);
    uint8_t *w = malloc(wrow * (size_t)m_sample), *y = malloc(yrow * (size_t)N);
    float *x = malloc(sizeof(float) * (size_t)(K * N)), *dst = malloc(sizeof(float) * (size_t)(m_sample * N));
    if (!w || !y || !x || !dst) { free(w); free(y); free(x); free(dst); return -1; }
    uint64_t s = 0x9E3779B97F4A7C15ull;
    fill_random_blocks(w, wrow * (size_t)m_sample, 0, &s);
    for (int64_t j = 0; j < K * N; ++j) x[j] = (float)((j * 2654435761u) % 2001) * 1e-3f - 1.0f;
    double best = 1e30, total = 0;
    for (int it = 0; it < warmup + iters; ++it) {
        const double t0 = now_s();
        for (int64_t n = 0; n < N; ++n) tq8->from_float(x + n * K, y + (size_t)n * yrow, K);
        for (int64_t n = 0; n < N; ++n)
            for (int64_t m = 0; m < m_sample; ++m)
                tq4->vec_dot((int)K, dst + n * m_sample + m, 0, w + (size_t)m * wrow, 0, y + (size_t)n * yrow, 0, 1);
        const double dt = now_s() - t0;
        if (it >= warmup) { total += dt; if (dt < best) best = dt; }
    }
    if (secs) { secs[0] = best; secs[1] = iters > 0 ? total / iters : 0; }
    if (checksum) { double c = 0; for (int64_t j = 0; j < m_sample * N; ++j) c += dst[j]; *checksum = c; }
    free(w); free(y); free(x); free(dst);
    return 0;
}

int ref_mul_mat_set_q4k(int n_mats, const int64_t *Ms, const int64_t *Ks, int64_t N, int n_threads,
                        int warmup, int iters, double *secs, double *checksum) {
    return ref_mul_mat_set_typed(n_mats, Ms, Ks, NULL, N, n_threads, warmup, iters, secs, checksum);
}

/* bench.py's `e2e_plugin` leg: a whole set of mul_mats (the Q4_K_M decode token) as ONE ggml graph on the backend plugin's first
 * device, driven the way a ggml host drives a backend — per step: ggml_backend_tensor_set_async of the step's activations from
 * (pinned, if the device offers a host buffer type) host memory, ggml_backend_graph_compute_async (or graph_plan_compute when
 * use_plan != 0 and the backend implements plans), ggml_backend_tensor_get_async of the results, ggml_backend_synchronize.
 * Wall-clock per step.  Inputs are 2-D views of ONE packed F32 tensor and every mul_mat's dst is placed (view_src / view_offs)
 * inside ONE packed result tensor, so a step moves one block up and one block down — the same traffic pattern as
 * q4k_plan_run_host.  input_ids[i]: mul_mats with equal ids read the same activations (q/k/v, gate/up).
 * Returns 0 or: -10 load failed, -11 no device, -12 supports_op says no, -13 init failed. */
int ref_plugin_graph_set(const char *path, int n_mats, const int64_t *Ms, const int64_t *Ks, const int *q6, const int *input_ids,
                         int64_t N, int use_plan, int warmup, int iters, double *secs, double *checksum,
                         int64_t *h2d_bytes, int64_t *d2h_bytes, int *pinned) {
    static ggml_backend_reg_t reg = NULL;
    if (!reg) reg = ggml_backend_load(path);
    if (!reg) return -10;
    if (ggml_backend_reg_dev_count(reg) < 1) return -11;
    ggml_backend_dev_t dev = ggml_backend_reg_dev_get(reg, 0);
    ggml_backend_t be = ggml_backend_dev_init(dev, NULL);
    if (!be) return -13;
    int rc = 0;
    struct ggml_init_params ipw = { ggml_tensor_overhead() * (size_t)(n_mats + 8), NULL, true };
    struct ggml_init_params ip = { ggml_tensor_overhead() * (size_t)(3 * n_mats + 16) +
                                   ggml_graph_overhead_custom((size_t)(4 * n_mats + 16), false), NULL, true };
    struct ggml_context *wctx = ggml_init(ipw), *ctx = ggml_init(ip);
    struct ggml_tensor **ws = calloc((size_t)n_mats, sizeof *ws), **ys = calloc((size_t)n_mats, sizeof *ys);
    struct ggml_tensor **xv = calloc((size_t)n_mats, sizeof *xv);
    int64_t *in_off = calloc((size_t)n_mats, sizeof *in_off);
    ggml_backend_buffer_t wbuf = NULL, buf = NULL, hbuf = NULL;
    ggml_backend_graph_plan_t plan = NULL;
    float *host_in = NULL, *host_out = NULL;
    if (!wctx || !ctx || !ws || !ys || !xv || !in_off) { rc = -1; goto out; }

    /* packed layout: one [N][K] block per distinct input id, one [N][M] block per mul_mat, blocks 256-byte aligned */
    int64_t in_floats = 0, out_floats = 0;
    for (int i = 0; i < n_mats; ++i) {
        int first = i;
        for (int j = 0; j < i; ++j) if (input_ids[j] == input_ids[i]) { first = j; break; }
        if (first == i) { in_off[i] = in_floats; in_floats += (N * Ks[i] + 63) / 64 * 64; }
        else in_off[i] = -1 - first;
    }
    for (int i = 0; i < n_mats; ++i) out_floats += (N * Ms[i] + 63) / 64 * 64;
    struct ggml_tensor *X = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, in_floats);
    struct ggml_tensor *Y = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, out_floats);
    struct ggml_cgraph *gf = ggml_new_graph_custom(ctx, (size_t)(4 * n_mats + 16), false);
    int64_t yo = 0;
    for (int i = 0; i < n_mats; ++i) {
        ws[i] = ggml_new_tensor_2d(wctx, q6 && q6[i] ? GGML_TYPE_Q6_K : GGML_TYPE_Q4_K, Ks[i], Ms[i]);
        xv[i] = in_off[i] >= 0 ? ggml_view_2d(ctx, X, Ks[i], N, (size_t)Ks[i] * 4, (size_t)in_off[i] * 4) : xv[-1 - in_off[i]];
        ys[i] = ggml_mul_mat(ctx, ws[i], xv[i]);
        ys[i]->view_src = Y;                               /* dst placed inside the packed result tensor */
        ys[i]->view_offs = (size_t)yo * 4;
        yo += (N * Ms[i] + 63) / 64 * 64;
        ggml_build_forward_expand(gf, ys[i]);
    }
    wbuf = ggml_backend_alloc_ctx_tensors(wctx, be);
    buf = ggml_backend_alloc_ctx_tensors(ctx, be);
    if (!wbuf || !buf) { rc = -2; goto out; }
    for (int i = 0; i < n_mats; ++i) if (!ggml_backend_dev_supports_op(dev, ys[i])) { rc = -12; goto out; }
    {   /* weights: random valid blocks, uploaded through the buffer's set_tensor */
        uint64_t s = 0x9E3779B97F4A7C15ull;
        for (int i = 0; i < n_mats; ++i) {
            const size_t nbytes = ggml_nbytes(ws[i]);
            uint8_t *tmp = malloc(nbytes);
            if (!tmp) { rc = -1; goto out; }
            fill_random_blocks(tmp, nbytes, q6 && q6[i], &s);
            ggml_backend_tensor_set(ws[i], tmp, 0, nbytes);
            free(tmp);
        }
    }
    {   /* host staging: the device's host (pinned) buffer type when it has one */
        const size_t hb = (size_t)(in_floats + out_floats) * 4;
        ggml_backend_buffer_type_t hbt = ggml_backend_dev_host_buffer_type(dev);
        if (hbt) hbuf = ggml_backend_buft_alloc_buffer(hbt, hb);
        if (pinned) *pinned = hbuf != NULL;
        host_in = hbuf ? (float *)ggml_backend_buffer_get_base(hbuf) : malloc(hb);
        if (!host_in) { rc = -1; goto out; }
        host_out = host_in + in_floats;
        for (int64_t j = 0; j < in_floats; ++j) host_in[j] = (float)((j * 2654435761u) % 2001) * 1e-3f - 1.0f;
        memset(host_out, 0, (size_t)out_floats * 4);
    }
    if (use_plan) plan = ggml_backend_graph_plan_create(be, gf);       /* NULL-safe below: backends without plans */
    double best = 1e30, total = 0;
    for (int it = 0; it < warmup + iters; ++it) {
        const double t0 = now_s();
        ggml_backend_tensor_set_async(be, X, host_in, 0, (size_t)in_floats * 4);
        const enum ggml_status st = plan ? ggml_backend_graph_plan_compute(be, plan) : ggml_backend_graph_compute_async(be, gf);
        if (st != GGML_STATUS_SUCCESS) { rc = -3; break; }
        ggml_backend_tensor_get_async(be, Y, host_out, 0, (size_t)out_floats * 4);
        ggml_backend_synchronize(be);
        const double dt = now_s() - t0;
        if (it >= warmup) { total += dt; if (dt < best) best = dt; }
    }
    if (secs) { secs[0] = best; secs[1] = iters > 0 ? total / iters : 0; }
    if (checksum) { double c = 0; for (int64_t j = 0; j < out_floats; ++j) c += host_out[j]; *checksum = c; }
    if (h2d_bytes) *h2d_bytes = in_floats * 4;
    if (d2h_bytes) *d2h_bytes = out_floats * 4;
out:
    if (plan) ggml_backend_graph_plan_free(be, plan);
    if (hbuf) ggml_backend_buffer_free(hbuf); else free(host_in);
    if (buf) ggml_backend_buffer_free(buf);
    if (wbuf) ggml_backend_buffer_free(wbuf);
    if (ctx) ggml_free(ctx);
    if (wctx) ggml_free(wctx);
    free(ws); free(ys); free(xv); free(in_off);
    ggml_backend_free(be);
    return rc;
}

/* The cheap F32 ops between the mul_mats of a decoder layer on ggml_backend_cpu, as the reference's llama graph strings them
 * together (build_norm / build_ffn): sum = add(x, res); y = mul(rms_norm(sum, eps), w); glu = swiglu_split(g, u).
 * x, res, g, u: [nrows][ne0] floats, w: [ne0].  Pins oracle/port.py's restatements (and the GPU kernels behind them). */
int ref_cheap_ops(const float *x, const float *res, const float *w, const float *g, const float *u, int64_t ne0, int64_t nrows, float eps,
                  int n_threads, float *y_sum, float *y_norm, float *y_glu) {
    ggml_backend_t cpu = ggml_backend_cpu_init();
    if (!cpu) return -1;
    ggml_backend_cpu_set_n_threads(cpu, n_threads);
    struct ggml_init_params ip = { ggml_tensor_overhead() * 16 + ggml_graph_overhead(), NULL, true };
    struct ggml_context *ctx = ggml_init(ip);
    if (!ctx) { ggml_backend_free(cpu); return -1; }
    struct ggml_tensor *tx = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, ne0, nrows), *tr = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, ne0, nrows);
    struct ggml_tensor *tw = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, ne0);
    struct ggml_tensor *tg = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, ne0, nrows), *tu = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, ne0, nrows);
    struct ggml_tensor *sum = ggml_add(ctx, tx, tr);
    struct ggml_tensor *nrm = ggml_mul(ctx, ggml_rms_norm(ctx, sum, eps), tw);
    struct ggml_tensor *glu = ggml_swiglu_split(ctx, tg, tu);
    struct ggml_cgraph *gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, nrm);
    ggml_build_forward_expand(gf, glu);
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, cpu);
    int rc = -2;
    if (buf) {
        const size_t nbytes = (size_t)ne0 * (size_t)nrows * sizeof(float);
        ggml_backend_tensor_set(tx, x, 0, nbytes); ggml_backend_tensor_set(tr, res, 0, nbytes);
        ggml_backend_tensor_set(tw, w, 0, (size_t)ne0 * sizeof(float));
        ggml_backend_tensor_set(tg, g, 0, nbytes); ggml_backend_tensor_set(tu, u, 0, nbytes);
        rc = ggml_backend_graph_compute(cpu, gf) == GGML_STATUS_SUCCESS ? 0 : -3;
        if (rc == 0) {
            ggml_backend_tensor_get(sum, y_sum, 0, nbytes); ggml_backend_tensor_get(nrm, y_norm, 0, nbytes);
            ggml_backend_tensor_get(glu, y_glu, 0, nbytes);
        }
        ggml_backend_buffer_free(buf);
    }
    ggml_free(ctx);
    ggml_backend_free(cpu);
    return rc;
}

/* ggml_rope_ext on ggml_backend_cpu: x [n_tokens][n_heads][ne0] floats, pos [n_tokens]; mode 0 = NORMAL, 2 = NEOX. */
int ref_rope(const float *x, const int32_t *pos, int64_t ne0, int64_t n_heads, int64_t n_tokens, int n_dims, int mode, int n_ctx_orig,
             float freq_base, float freq_scale, float ext_factor, float attn_factor, float beta_fast, float beta_slow, int n_threads, float *y) {
    ggml_backend_t cpu = ggml_backend_cpu_init();
    if (!cpu) return -1;
    ggml_backend_cpu_set_n_threads(cpu, n_threads);
    struct ggml_init_params ip = { ggml_tensor_overhead() * 8 + ggml_graph_overhead(), NULL, true };
    struct ggml_context *ctx = ggml_init(ip);
    if (!ctx) { ggml_backend_free(cpu); return -1; }
    struct ggml_tensor *tx = ggml_new_tensor_3d(ctx, GGML_TYPE_F32, ne0, n_heads, n_tokens);
    struct ggml_tensor *tp = ggml_new_tensor_1d(ctx, GGML_TYPE_I32, n_tokens);
    struct ggml_tensor *ty = ggml_rope_ext(ctx, tx, tp, NULL, n_dims, mode, n_ctx_orig, freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow);
    struct ggml_cgraph *gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, ty);
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, cpu);
    int rc = -2;
    if (buf) {
        ggml_backend_tensor_set(tx, x, 0, ggml_nbytes(tx));
        ggml_backend_tensor_set(tp, pos, 0, ggml_nbytes(tp));
        rc = ggml_backend_graph_compute(cpu, gf) == GGML_STATUS_SUCCESS ? 0 : -3;
        if (rc == 0) ggml_backend_tensor_get(ty, y, 0, ggml_nbytes(ty));
        ggml_backend_buffer_free(buf);
    }
    ggml_free(ctx);
    ggml_backend_free(cpu);
    return rc;
}

/* --split-mode row through the plugin: the weights in the buffer type the registry's "ggml_backend_split_buffer_type" proc
 * returns (L/src/llama-model.cpp:375-401), activations and result in the main device's default buffer, one ggml_mul_mat node
 * computed by the main device's backend.  w_back (optional) receives the weights read back through get_tensor (the gather).
 * q6k: 0 = Q4_K weights, 1 = Q6_K.  Returns 0, or -20 if the plugin has no such proc, -21 no buffer type, -12 not supported. */
typedef ggml_backend_buffer_type_t (*split_buft_fn)(int main_device, const float *tensor_split);
int ref_plugin_split_mul_mat(const char *path, int q6k, const void *w, int64_t M, int64_t K, const float *x, int64_t N, float *dst,
                             const float *tensor_split, void *w_back, int *n_devices) {
    static ggml_backend_reg_t reg = NULL;
    if (!reg) reg = ggml_backend_load(path);
    if (!reg) return -10;
    if (n_devices) *n_devices = (int)ggml_backend_reg_dev_count(reg);
    if (ggml_backend_reg_dev_count(reg) < 1) return -11;
    split_buft_fn fn = (split_buft_fn)ggml_backend_reg_get_proc_address(reg, "ggml_backend_split_buffer_type");
    if (!fn) return -20;
    ggml_backend_buffer_type_t sbuft = fn(0, tensor_split);
    if (!sbuft) return -21;
    ggml_backend_dev_t dev = ggml_backend_reg_dev_get(reg, 0);
    ggml_backend_t be = ggml_backend_dev_init(dev, NULL);
    if (!be) return -13;
    struct ggml_init_params ip = { ggml_tensor_overhead() * 8 + ggml_graph_overhead(), NULL, true };
    struct ggml_context *cw = ggml_init(ip), *cc = ggml_init(ip);
    struct ggml_tensor *tw = ggml_new_tensor_2d(cw, q6k ? GGML_TYPE_Q6_K : GGML_TYPE_Q4_K, K, M);
    ggml_backend_buffer_t bw = ggml_backend_alloc_ctx_tensors_from_buft(cw, sbuft);
    struct ggml_tensor *tx = ggml_new_tensor_2d(cc, GGML_TYPE_F32, K, N);
    struct ggml_tensor *ty = ggml_mul_mat(cc, tw, tx);
    struct ggml_cgraph *gf = ggml_new_graph(cc);
    ggml_build_forward_expand(gf, ty);
    ggml_backend_buffer_t bc = ggml_backend_alloc_ctx_tensors(cc, be);
    int rc = (bw && bc) ? 0 : -2;
    if (rc == 0 && !(ggml_backend_dev_supports_buft(dev, sbuft) && ggml_backend_dev_supports_op(dev, ty))) rc = -12;
    if (rc == 0) {
        ggml_backend_tensor_set(tw, w, 0, ggml_nbytes(tw));
        ggml_backend_tensor_set(tx, x, 0, ggml_nbytes(tx));
        for (int it = 0; it < 2 && rc == 0; ++it)          /* twice: the second compute reuses streams, events and workspaces */
            if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) rc = -3;
        if (rc == 0) {
            ggml_backend_synchronize(be);
            ggml_backend_tensor_get(ty, dst, 0, ggml_nbytes(ty));
            if (w_back) ggml_backend_tensor_get(tw, w_back, 0, ggml_nbytes(tw));
        }
    }
    if (bc) ggml_backend_buffer_free(bc);
    if (bw) ggml_backend_buffer_free(bw);
    ggml_free(cc);
    ggml_free(cw);
    ggml_backend_free(be);
    return rc;
}
