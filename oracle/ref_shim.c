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
int ref_mul_mat_set_typed(int n_mats, const int64_t *Ms, const int64_t *Ks, const int *q6, int64_t N, int n_threads,
                          int warmup, int iters, double *secs, double *checksum) {
    ggml_backend_t cpu = ggml_backend_cpu_init();
    if (!cpu) return -1;
    ggml_backend_cpu_set_n_threads(cpu, n_threads);
    struct ggml_init_params ip = { ggml_tensor_overhead() * (size_t)(3 * n_mats + 8) +
                                   ggml_graph_overhead_custom((size_t)(4 * n_mats + 16), false), NULL, true };
    struct ggml_context *ctx = ggml_init(ip);
    if (!ctx) { ggml_backend_free(cpu); return -1; }
    struct ggml_cgraph *gf = ggml_new_graph_custom(ctx, (size_t)(4 * n_mats + 16), false);
    struct ggml_tensor **ws = calloc((size_t)n_mats, sizeof *ws), **xs = calloc((size_t)n_mats, sizeof *xs),
                       **ys = calloc((size_t)n_mats, sizeof *ys);
    for (int i = 0; i < n_mats; ++i) {
        ws[i] = ggml_new_tensor_2d(ctx, q6 && q6[i] ? GGML_TYPE_Q6_K : GGML_TYPE_Q4_K, Ks[i], Ms[i]);
        xs[i] = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, Ks[i], N);
        ys[i] = ggml_mul_mat(ctx, ws[i], xs[i]);
        ggml_build_forward_expand(gf, ys[i]);
    }
    int rc = 0;
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors_from_buft(ctx, ggml_backend_cpu_buffer_type());
    if (!buf) { rc = -2; goto done; }
    uint64_t s = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < n_mats; ++i) {
        uint8_t *p = (uint8_t *)ws[i]->data;
        const size_t nbytes = ggml_nbytes(ws[i]);
        for (size_t o = 0; o + 8 <= nbytes; o += 8) {
            s ^= s << 13; s ^= s >> 7; s ^= s << 17;
            memcpy(p + o, &s, 8);
        }
        if (q6 && q6[i]) { for (size_t o = 0; o < nbytes; o += 210) { p[o + 208] = 0x00; p[o + 209] = 0x1c; } }
        else { for (size_t o = 0; o < nbytes; o += 144) { p[o] = 0x00; p[o + 1] = 0x2c; p[o + 2] = 0x00; p[o + 3] = 0x28; } }
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
    ggml_free(ctx);
    free(ws); free(xs); free(ys);
    ggml_backend_free(cpu);
    return rc;
}

int ref_mul_mat_set_q4k(int n_mats, const int64_t *Ms, const int64_t *Ks, int64_t N, int n_threads,
                        int warmup, int iters, double *secs, double *checksum) {
    return ref_mul_mat_set_typed(n_mats, Ms, Ks, NULL, N, n_threads, warmup, iters, secs, checksum);
}
