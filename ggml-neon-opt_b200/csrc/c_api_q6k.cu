// c_api_q6k.cu — the Q6_K entries of include/q4k_b200.h over gemv_q6k.cu.
// Argument validation, column tiling, error mapping.  No CPU fallback: if the kernel cannot run, the call fails.
#include "../../include/q4k_b200.h"

#include <cstdlib>

#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {
int note_cuda_error(cudaError_t e);
}
using namespace q4k;

namespace {


bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int check_w(const void *w, int64_t nb, int64_t stride) {
    if (stride < nb * Q4K_BLOCK_Q6K || (stride & 1) || (reinterpret_cast<uintptr_t>(w) & 1u)) return Q4K_ERR_BAD_SHAPE;
    return Q4K_OK;
}

// columns [0, N) in passes of at most 8, fewer when K is so long that 8 columns of Q8_K do not fit in shared memory
int run_columns(const void *w, int64_t M, int64_t K, int64_t w_stride, const void *xq, const float *xf, int64_t N,
                int64_t x_stride, float *dst, int64_t dst_stride, int32_t *dump, cudaStream_t st) {
    int64_t step = Q4K_GEMV_MAX_N;
    const int64_t nb = K / Q4K_QK_K;
    for (int64_t n0 = 0; n0 < N;) {
        const int nn = (int)((N - n0) < step ? (N - n0) : step);
        const cudaError_t e = launch_gemv_q6k(w, M, K, w_stride,
                                              xq ? (const char *)xq + n0 * x_stride : nullptr,
                                              xf ? (const float *)((const char *)xf + n0 * x_stride) : nullptr, nn, x_stride,
                                              (float *)((char *)dst + n0 * dst_stride), dst_stride,
                                              dump ? dump + n0 * M * nb : nullptr, knobs().pdl != 0, st);
        if (e == cudaErrorNotSupported && nn > 1) { step = nn / 2; continue; }
        if (e == cudaErrorNotSupported) return Q4K_ERR_UNSUPPORTED;
        if (e != cudaSuccess) return note_cuda_error(e);
        n0 += nn;
    }
    return Q4K_OK;
}

int zero_dst(float *dst, int64_t M, int64_t N, int64_t dst_stride, cudaStream_t st) {
    for (int64_t n = 0; n < N; ++n) {
        const cudaError_t e = cudaMemsetAsync((char *)dst + n * dst_stride, 0, (size_t)M * 4, st);
        if (e != cudaSuccess) return note_cuda_error(e);
    }
    return Q4K_OK;
}

}  // namespace

extern "C" {

size_t q6k_mul_mat_workspace_bytes(int64_t K, int64_t N) {
    if (K <= 0 || N <= 0 || K % Q4K_QK_K != 0) return 0;
    size_t need = (size_t)N * (size_t)(K / Q4K_QK_K) * Q4K_BLOCK_Q8K;
    const int gmin = knobs().q6k_gemm_min;
    if (gmin > 0 && N >= gmin) need = need > gemm_q6k_workspace_bytes(K, N) ? need : gemm_q6k_workspace_bytes(K, N);
    return need;
}

int q6k_mul_mat_q8(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                   const void *xq_dev, int64_t N, int64_t xq_row_stride_bytes,
                   float *dst_dev, int64_t dst_col_stride_bytes, void *stream) {
    if (M < 0 || N < 0 || K < 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (M == 0 || N == 0) return Q4K_OK;
    if (!dst_dev) return Q4K_ERR_BAD_POINTER;
    if (dst_col_stride_bytes < M * 4 || (dst_col_stride_bytes & 3)) return Q4K_ERR_BAD_SHAPE;
    if (K == 0) return zero_dst(dst_dev, M, N, dst_col_stride_bytes, (cudaStream_t)stream);
    if (!w_dev || !xq_dev) return Q4K_ERR_BAD_POINTER;
    const int64_t nb = K / Q4K_QK_K;
    if (check_w(w_dev, nb, w_row_stride_bytes)) return Q4K_ERR_BAD_SHAPE;
    if (xq_row_stride_bytes < nb * Q4K_BLOCK_Q8K || (xq_row_stride_bytes & 3) || (reinterpret_cast<uintptr_t>(xq_dev) & 3u)) return Q4K_ERR_BAD_SHAPE;
    return run_columns(w_dev, M, K, w_row_stride_bytes, xq_dev, nullptr, N, xq_row_stride_bytes, dst_dev, dst_col_stride_bytes,
                       nullptr, (cudaStream_t)stream);
}

int q6k_mul_mat(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                float *dst_dev, int64_t dst_col_stride_bytes,
                void *workspace_dev, size_t workspace_bytes, void *stream) {
    if (M < 0 || N < 0 || K < 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (M == 0 || N == 0) return Q4K_OK;
    if (!dst_dev) return Q4K_ERR_BAD_POINTER;
    if (dst_col_stride_bytes < M * 4 || (dst_col_stride_bytes & 3)) return Q4K_ERR_BAD_SHAPE;
    cudaStream_t st = (cudaStream_t)stream;
    if (K == 0) return zero_dst(dst_dev, M, N, dst_col_stride_bytes, st);
    if (!w_dev || !x_dev) return Q4K_ERR_BAD_POINTER;
    const int64_t nb = K / Q4K_QK_K;
    if (check_w(w_dev, nb, w_row_stride_bytes)) return Q4K_ERR_BAD_SHAPE;
    if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
    const int64_t fuse_limit = knobs().q6k_fuse_limit;                          // default 32768 (tools/fuse_sweep.py)
    if (N <= Q4K_GEMV_MAX_N && N * K <= fuse_limit) {             // decode: quantizer fused into the launch
        const int rc = run_columns(w_dev, M, K, w_row_stride_bytes, nullptr, x_dev, N, x_row_stride_bytes, dst_dev,
                                   dst_col_stride_bytes, nullptr, st);
        if (rc != Q4K_ERR_UNSUPPORTED) return rc;
    }
    if (workspace_bytes < q6k_mul_mat_workspace_bytes(K, N)) return Q4K_ERR_WORKSPACE;
    if (!workspace_dev || (reinterpret_cast<uintptr_t>(workspace_dev) & 3u)) return Q4K_ERR_BAD_POINTER;
    if (knobs().q6k_gemm_min > 0 && N >= knobs().q6k_gemm_min) {                 // prefill: tcgen05 kind::f16 GEMM
        if (reinterpret_cast<uintptr_t>(workspace_dev) & 15u) return Q4K_ERR_BAD_POINTER;
        const cudaError_t eg = launch_gemm_q6k_f32(w_dev, M, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_dev,
                                                   dst_col_stride_bytes, workspace_dev, st);
        if (eg != cudaErrorNotSupported) return eg == cudaSuccess ? Q4K_OK : note_cuda_error(eg);
    }
    const cudaError_t e = launch_quantize_q8_K(x_dev, workspace_dev, K, N, x_row_stride_bytes, st);
    if (e != cudaSuccess) return note_cuda_error(e);
    return run_columns(w_dev, M, K, w_row_stride_bytes, workspace_dev, nullptr, N, nb * Q4K_BLOCK_Q8K, dst_dev,
                       dst_col_stride_bytes, nullptr, st);
}

/* parity tooling: the prefill GEMM's DUMP instantiation at any N — sumi[(n*M + m)*nb + b] = the integers its epilogue read from
 * tensor memory, dst = the product floats */
int q6k_mul_mat_dump(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x_dev, int64_t N,
                     int64_t x_row_stride_bytes, int32_t *sumi_dev, float *dst_dev, void *workspace_dev, size_t workspace_bytes,
                     void *stream) {
    if (M <= 0 || N <= 0 || K <= 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (!w_dev || !x_dev || !sumi_dev || !dst_dev || !workspace_dev) return Q4K_ERR_BAD_POINTER;
    if (check_w(w_dev, K / Q4K_QK_K, w_row_stride_bytes)) return Q4K_ERR_BAD_SHAPE;
    if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev) || !aligned16(workspace_dev)) return Q4K_ERR_BAD_SHAPE;
    if (workspace_bytes < gemm_q6k_workspace_bytes(K, N)) return Q4K_ERR_WORKSPACE;
    const cudaError_t e = launch_gemm_q6k_f32(w_dev, M, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_dev, M * 4,
                                              workspace_dev, (cudaStream_t)stream, sumi_dev);
    return e == cudaSuccess ? Q4K_OK : e == cudaErrorNotSupported ? Q4K_ERR_UNSUPPORTED : note_cuda_error(e);
}

size_t q6k_gemm_workspace_bytes(int64_t K, int64_t N) {
    if (K <= 0 || N <= 0 || K % Q4K_QK_K != 0) return 0;
    return gemm_q6k_workspace_bytes(K, N);
}

int q6k_debug_partials(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                       const void *xq_dev, int64_t N, int64_t xq_row_stride_bytes,
                       int32_t *sumi_dev, float *dst_dev, void *stream) {
    if (M <= 0 || N <= 0 || K <= 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (!w_dev || !xq_dev || !sumi_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
    const int64_t nb = K / Q4K_QK_K;
    if (check_w(w_dev, nb, w_row_stride_bytes)) return Q4K_ERR_BAD_SHAPE;
    if (xq_row_stride_bytes < nb * Q4K_BLOCK_Q8K || (xq_row_stride_bytes & 3)) return Q4K_ERR_BAD_SHAPE;
    return run_columns(w_dev, M, K, w_row_stride_bytes, xq_dev, nullptr, N, xq_row_stride_bytes, dst_dev, M * 4, sumi_dev,
                       (cudaStream_t)stream);
}

}  // extern "C"
