// c_api.cu — the C ABI of include/q4k_b200.h over the kernels of this directory.
// Argument validation, dispatch (GEMV for N <= 8, tcgen05 GEMM above), error mapping.  No CPU
// fallback of any kind: if a kernel cannot run, the call fails with a negative code.
#include "../../include/q4k_b200.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "knobs.h"
#include "q4k_common.cuh"
#include "q4k_kernels.h"

#define Q4K_STR2(x) #x
#define Q4K_STR(x) Q4K_STR2(x)

namespace q4k {

static std::atomic<int64_t> g_launches{0};
static thread_local int t_last_cuda_error = 0;

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
void add_launches(int64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int64_t launch_count() { return g_launches.load(std::memory_order_relaxed); }
bool gemv_exact_policy(int N);

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static int cuda_fail(cudaError_t e) {
    t_last_cuda_error = (int)e;
    return Q4K_ERR_CUDA;
}
int note_cuda_error(cudaError_t e) { return cuda_fail(e); }

static bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Float order of the decode path (N <= 8).  Default: the tensor-pipe kernel (gemv_q4k_mma.cu), exact integers and a fixed
// launch-independent float association, within 1e-5 of the reference.  Q4K_GEMV_ORDER=exact (or
// q4k_debug_set_gemv_order(1)) selects the kernel that reproduces the reference's AVX2 accumulation order bit for bit
// (gemv_q4k_tma.cu) at ~10-25% less throughput on small matrices.
constexpr int64_t Q4K_GEMV_GEMM_MIN_N = 12;   // largest N served by passes of the decode kernel when the GEMM is available
bool gemv_exact_policy(int N);
static bool gemv_exact_for(int N) { return gemv_exact_policy(N); }
bool gemv_exact_policy(int N) {
    (void)N;
    return knobs().order_mode == 1;
}

}  // namespace q4k

using namespace q4k;

extern "C" {

int q4k_abi_version(void) { return 1; }

const char *q4k_build_info(void) {
    return "q4k_b200 sm_100a, CUDA " Q4K_STR(CUDART_VERSION) ", built " __DATE__;
}

int q4k_device_ok(int device) {
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device) != cudaSuccess) return 0;
    return major == 10;
}

int q4k_last_cuda_error(void) { return t_last_cuda_error; }

int64_t q4k_kernel_launches(void) { return g_launches.load(std::memory_order_relaxed); }

int q4k_quantize_q8_K(const float *x_dev, void *y_dev, int64_t k, int64_t nrows, int64_t x_row_stride_bytes, void *stream) {
    if (k < 0 || nrows < 0 || k % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (k == 0 || nrows == 0) return Q4K_OK;
    if (!x_dev || !y_dev) return Q4K_ERR_BAD_POINTER;
    if (x_row_stride_bytes < k * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
    if (reinterpret_cast<uintptr_t>(y_dev) & 3u) return Q4K_ERR_BAD_POINTER;
    cudaError_t e = launch_quantize_q8_K(x_dev, y_dev, k, nrows, x_row_stride_bytes, (cudaStream_t)stream);
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

size_t q4k_mul_mat_workspace_bytes(int64_t K, int64_t N) {
    if (K <= 0 || N <= 0 || K % Q4K_QK_K != 0) return 0;
    const size_t nb = (size_t)(K / Q4K_QK_K);
    size_t need = (size_t)N * nb * Q4K_BLOCK_Q8K;                        // block_q8_K rows (decode path)
    if (N > Q4K_GEMV_MAX_N) {                                            // prefill path: int8 + mins tiles of whole 128-column tiles, scales
        const size_t gemm = q4k::gemm_workspace_bytes(K, N);
        if (gemm > need) need = gemm;

    }
    if (N == 1 && q4k::n1s_workspace_bytes(K) > need) need = q4k::n1s_workspace_bytes(K);      // streamed single-column kernel: whole rounds of staged blocks
    if (N >= knobs().wide_min && N <= knobs().midn_max) {                // batched decode kernel: staged Q8_K, padded to 8 columns
        const size_t wide = q4k::wide_workspace_bytes(K, N);
        if (wide > need) need = wide;
    }
    return need;
}

int q4k_mul_mat_q8(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                   const void *xq_dev, int64_t N, int64_t xq_row_stride_bytes,
                   float *dst_dev, int64_t dst_col_stride_bytes, void *stream) {
    if (M < 0 || N < 0 || K < 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (M == 0 || N == 0) return Q4K_OK;
    if (!dst_dev) return Q4K_ERR_BAD_POINTER;
    const int64_t nb = K / Q4K_QK_K;
    if (K == 0) {        // empty contraction: the reference's vec_dot loop runs zero times -> 0.0f
        for (int64_t n = 0; n < N; ++n) {
            cudaError_t e = cudaMemsetAsync((char *)dst_dev + n * dst_col_stride_bytes, 0, (size_t)M * 4, (cudaStream_t)stream);
            if (e != cudaSuccess) return cuda_fail(e);
        }
        return Q4K_OK;
    }
    if (!w_dev || !xq_dev) return Q4K_ERR_BAD_POINTER;
    if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
    if (xq_row_stride_bytes < nb * Q4K_BLOCK_Q8K || (xq_row_stride_bytes & 3) || (reinterpret_cast<uintptr_t>(xq_dev) & 3u)) return Q4K_ERR_BAD_SHAPE;
    if (dst_col_stride_bytes < M * 4 || (dst_col_stride_bytes & 3)) return Q4K_ERR_BAD_SHAPE;
    cudaStream_t st = (cudaStream_t)stream;
    if (N > Q4K_GEMV_MAX_N) {
        cudaError_t e = launch_gemm_q4k_q8k(w_dev, M, K, w_row_stride_bytes, xq_dev, N, xq_row_stride_bytes, dst_dev, dst_col_stride_bytes, st);
        if (e == cudaSuccess) return Q4K_OK;
        if (e != cudaErrorNotSupported) return cuda_fail(e);
        // shape the tensor path does not take (yet): same GPU arithmetic, eight columns per pass
    }
    for (int64_t n0 = 0; n0 < N; n0 += Q4K_GEMV_MAX_N) {
        const int nn = (int)((N - n0) < Q4K_GEMV_MAX_N ? (N - n0) : Q4K_GEMV_MAX_N);
        const bool exact = N <= Q4K_GEMV_MAX_N ? gemv_exact_for(nn) : false;
        cudaError_t e = launch_gemv_q4k_q8k(w_dev, M, K, w_row_stride_bytes,
                                            (const char *)xq_dev + n0 * xq_row_stride_bytes, nn, xq_row_stride_bytes,
                                            (float *)((char *)dst_dev + n0 * dst_col_stride_bytes), dst_col_stride_bytes, exact, st);
        if (e != cudaSuccess) return cuda_fail(e);
    }
    return Q4K_OK;
}

int q4k_mul_mat(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                float *dst_dev, int64_t dst_col_stride_bytes,
                void *workspace_dev, size_t workspace_bytes, void *stream) {
    if (M < 0 || N < 0 || K < 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (M == 0 || N == 0) return Q4K_OK;
    if (K > 0) {
        if (workspace_bytes < q4k_mul_mat_workspace_bytes(K, N)) return Q4K_ERR_WORKSPACE;
        if (!workspace_dev) return Q4K_ERR_BAD_POINTER;
        // (the bit-exact AVX2 order, when asked for, exists only in the decode kernel)
        // (two columns of a short K stay on the decode kernel: 5632 x 2048 N = 2 4.5 vs 5.1 us; from K = 4096 the batched kernel wins)
        const bool wide = N >= knobs().wide_min && N <= knobs().midn_max && !(N <= Q4K_GEMV_MAX_N && gemv_exact_for((int)N)) &&
                          !(N == 2 && K < 4096);
        // one column of a matrix that streams tens of MB or has a long K: quantize once, stream the activations next to the weights
        // (gemv_q4k_n1s.cu; the bits of the decode kernel below)
        // Measured (profiles/r2_n1s_sweep.txt): wins where K is long and there are enough tiles for every SM — 4096 x 14336
        // 11.85 -> 9.45 us, 4096 x 28672 18.3 -> 14.5, 8192 x 28672 27.6 -> 26.7; loses on short K (14336 x 4096 9.1 -> 9.7) and on
        // few rows (1024 x 28672: 32 tile pairs for 148 SMs).
        if (N == 1 && !gemv_exact_for(1) && knobs().gemv_config == 0 && !knobs().mma_off &&
            ((K >= knobs().n1s_min_k && M >= 2048) ||
             (knobs().n1s_min_bytes > 0 && M * (K / Q4K_QK_K) * Q4K_BLOCK_Q4K >= knobs().n1s_min_bytes))) {
            if (!w_dev || !x_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
            const int64_t nb = K / Q4K_QK_K;
            if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
            if (!aligned16(x_dev) || !aligned16(workspace_dev)) return Q4K_ERR_BAD_SHAPE;
            const cudaError_t e = launch_gemv_q4k_n1s(w_dev, M, K, w_row_stride_bytes, x_dev, dst_dev, workspace_dev, knobs().pdl != 0,
                                                      (cudaStream_t)stream);
            if (e == cudaSuccess) return Q4K_OK;
            if (e != cudaErrorNotSupported) return cuda_fail(e);
        }
        // decode shapes: one launch — every CTA quantizes the (small) activations into its own shared memory
        if (!wide && N <= Q4K_GEMV_MAX_N && N * K <= knobs().fuse_limit && knobs().gemv_config == 0) {
            if (!w_dev || !x_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
            const int64_t nb = K / Q4K_QK_K;
            if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
            if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
            if (dst_col_stride_bytes < M * 4 || (dst_col_stride_bytes & 3)) return Q4K_ERR_BAD_SHAPE;
            const cudaError_t e = launch_gemv_q4k_q8k_tma(w_dev, M, K, w_row_stride_bytes, nullptr, x_dev, (int)N, x_row_stride_bytes,
                                                          dst_dev, dst_col_stride_bytes, gemv_exact_for((int)N), 0,
                                                          knobs().pdl != 0, (cudaStream_t)stream);
            if (e == cudaSuccess) return Q4K_OK;
            if (e != cudaErrorNotSupported) return cuda_fail(e);
        }
        // batches between the decode kernel and the GEMM (8 < N <= midn_max): the batched tensor-pipe decode kernel — the
        // weights stream once per 32 columns, the activations stream through shared memory next to them (gemv_q4k_wide.cu)
        if (wide) {
            if (!w_dev || !x_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
            const int64_t nb = K / Q4K_QK_K;
            if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
            if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
            if (dst_col_stride_bytes < M * 4 || (dst_col_stride_bytes & 3) || !aligned16(workspace_dev)) return Q4K_ERR_BAD_SHAPE;
            const cudaError_t e = launch_gemv_q4k_wide(w_dev, M, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_dev,
                                                       dst_col_stride_bytes, workspace_dev, knobs().pdl != 0, (cudaStream_t)stream);
            if (e == cudaSuccess) return Q4K_OK;
            if (e != cudaErrorNotSupported) return cuda_fail(e);
        }
        // prefill shapes: planar Q8 quantizer + tensor-core GEMM (tcgen05 kind::i8).  Its time does not depend on N up to a
        // 128-column tile, so 9..12 columns are cheaper as two passes of the decode kernel (the second pass finds the
        // weights in L2): measured at six shapes, profiles/r1_midn_sweep.txt (4096x14336 N = 9: 39 vs 49 us)
        if (N > Q4K_GEMV_GEMM_MIN_N && knobs().gemm_enabled) {
            if (!w_dev || !x_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
            const int64_t nb = K / Q4K_QK_K;
            if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
            if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
            if (dst_col_stride_bytes < M * 4 || (dst_col_stride_bytes & 3) || !aligned16(workspace_dev)) return Q4K_ERR_BAD_SHAPE;
            const cudaError_t e = launch_gemm_q4k_f32(w_dev, M, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_dev,
                                                      dst_col_stride_bytes, workspace_dev, (cudaStream_t)stream);
            return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
        }
        const int rc = q4k_quantize_q8_K(x_dev, workspace_dev, K, N, x_row_stride_bytes, stream);
        if (rc != Q4K_OK) return rc;
    }
    return q4k_mul_mat_q8(w_dev, M, K, w_row_stride_bytes, workspace_dev, N, (K / Q4K_QK_K) * Q4K_BLOCK_Q8K,
                          dst_dev, dst_col_stride_bytes, stream);
}

int q4k_mul_mat_peers_ws(const void *w_dev, int64_t M_local, int64_t K, int64_t w_row_stride_bytes,
                         const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                         float *const *dst_peers, int n_peers, int64_t dst_row0, int64_t dst_col_stride_bytes,
                         void *workspace_dev, size_t workspace_bytes, void *stream) {
    if (M_local < 0 || N < 0 || K <= 0 || K % Q4K_QK_K != 0 || dst_row0 < 0) return Q4K_ERR_BAD_SHAPE;
    if (N > Q4K_GEMV_MAX_N || n_peers < 1 || n_peers > 8) return Q4K_ERR_UNSUPPORTED;
    if (M_local == 0 || N == 0) return Q4K_OK;
    if (!w_dev || !x_dev || !dst_peers) return Q4K_ERR_BAD_POINTER;
    for (int g = 0; g < n_peers; ++g) if (!dst_peers[g]) return Q4K_ERR_BAD_POINTER;
    const int64_t nb = K / Q4K_QK_K;
    if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
    if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
    if (dst_col_stride_bytes < (dst_row0 + M_local) * 4 || (dst_col_stride_bytes & 3)) return Q4K_ERR_BAD_SHAPE;
    const void *xq = nullptr;
    int64_t xq_stride = x_row_stride_bytes;
    if (N * K > knobs().fuse_limit && workspace_dev && workspace_bytes >= (size_t)(N * nb * Q4K_BLOCK_Q8K)) {
        // long activations: quantize once in a kernel of its own instead of once per CTA (same rule as q4k_mul_mat)
        const cudaError_t eq = launch_quantize_q8_K(x_dev, workspace_dev, K, N, x_row_stride_bytes, (cudaStream_t)stream);
        if (eq != cudaSuccess) return cuda_fail(eq);
        xq = workspace_dev;
        xq_stride = nb * Q4K_BLOCK_Q8K;
    }
    const cudaError_t e = launch_gemv_q4k_q8k_tma(w_dev, M_local, K, w_row_stride_bytes, xq, xq ? nullptr : x_dev, (int)N, xq_stride,
                                                  nullptr, dst_col_stride_bytes, gemv_exact_for((int)N), 0, knobs().pdl != 0,
                                                  (cudaStream_t)stream, dst_peers, n_peers, dst_row0);
    if (e == cudaErrorNotSupported) return Q4K_ERR_UNSUPPORTED;     // activations too large for one shared-memory residency
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

size_t q4k_ll_buffer_bytes(int64_t M_full, int64_t n_max) {
    if (M_full <= 0 || n_max <= 0) return 0;
    return (size_t)2 * (size_t)n_max * (size_t)M_full * 8 + 2 * 8 * 8;       // data pairs + presence pairs
}

int q4k_mul_mat_peers_ll(const void *w_dev, int64_t M_local, int64_t K, int64_t w_row_stride_bytes,
                         const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                         void *const *ll_peers, int world, int rank, int64_t M_full, int64_t n_max, int64_t dst_row0,
                         float *dst_dev, int64_t dst_col_stride_bytes, void *cnt_dev,
                         void *workspace_dev, size_t workspace_bytes, void *stream) {
    if (M_local < 0 || N < 0 || K <= 0 || K % Q4K_QK_K != 0 || dst_row0 < 0 || M_full < dst_row0 + M_local) return Q4K_ERR_BAD_SHAPE;
    if (N > Q4K_GEMV_MAX_N || N > n_max || world < 1 || world > 8 || rank < 0 || rank >= world) return Q4K_ERR_UNSUPPORTED;
    if (N == 0) return Q4K_OK;
    if (!ll_peers || !dst_dev || !cnt_dev || (M_local > 0 && (!w_dev || !x_dev))) return Q4K_ERR_BAD_POINTER;
    if (dst_col_stride_bytes < M_full * 4 || (dst_col_stride_bytes & 3) || (reinterpret_cast<uintptr_t>(cnt_dev) & 7u)) return Q4K_ERR_BAD_SHAPE;
    LLExchange ll;
    for (int r = 0; r < 8; ++r) ll.peers[r] = nullptr;
    for (int r = 0; r < world; ++r) {
        if (!ll_peers[r] || (reinterpret_cast<uintptr_t>(ll_peers[r]) & 7u)) return Q4K_ERR_BAD_POINTER;
        ll.peers[r] = ll_peers[r];
    }
    ll.local = ll_peers[rank]; ll.half_pairs = n_max * M_full; ll.M_full = M_full;
    ll.cnt = static_cast<uint32_t *>(cnt_dev); ll.dst = dst_dev; ll.dst_stride = dst_col_stride_bytes / 4; ll.world = world; ll.rank = rank;
    cudaStream_t st = (cudaStream_t)stream;
    if (M_local == 0) {                        // nothing to compute here: gather everybody else's rows
        const cudaError_t e = launch_ll_gather(ll, (int)N, knobs().pdl != 0, st);
        return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
    }
    const int64_t nb = K / Q4K_QK_K;
    if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
    if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
    // one column of a long K with enough rows per rank: the streamed kernel, as q4k_mul_mat would choose for the shard alone
    if (N == 1 && !gemv_exact_for(1) && knobs().gemv_config == 0 && !knobs().mma_off && K >= knobs().n1s_min_k && M_local >= knobs().n1s_ll_min_rows &&
        workspace_dev && workspace_bytes >= n1s_workspace_bytes(K) && aligned16(workspace_dev)) {
        const cudaError_t e1 = launch_gemv_q4k_n1s(w_dev, M_local, K, w_row_stride_bytes, x_dev, nullptr, workspace_dev, knobs().pdl != 0, st, &ll, dst_row0);
        if (e1 != cudaErrorNotSupported) return e1 == cudaSuccess ? Q4K_OK : cuda_fail(e1);
    }
    const void *xq = nullptr;
    int64_t xq_stride = x_row_stride_bytes;
    if (N * K > knobs().fuse_limit && workspace_dev && workspace_bytes >= (size_t)(N * nb * Q4K_BLOCK_Q8K)) {
        const cudaError_t eq = launch_quantize_q8_K(x_dev, workspace_dev, K, N, x_row_stride_bytes, st);
        if (eq != cudaSuccess) return cuda_fail(eq);
        xq = workspace_dev;
        xq_stride = nb * Q4K_BLOCK_Q8K;
    }
    const cudaError_t e = launch_gemv_q4k_q8k_tma(w_dev, M_local, K, w_row_stride_bytes, xq, xq ? nullptr : x_dev, (int)N, xq_stride,
                                                  nullptr, 0, gemv_exact_for((int)N), 0, knobs().pdl != 0, st, nullptr, 0, dst_row0,
                                                  nullptr, 0, &ll);
    if (e == cudaErrorNotSupported) return Q4K_ERR_UNSUPPORTED;     // activations too large for one shared-memory residency
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

int q4k_mul_mat_reuse(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes, int64_t N,
                      float *dst_dev, int64_t dst_col_stride_bytes, void *workspace_dev, size_t workspace_bytes, void *stream) {
    if (M < 0 || N < 0 || K <= 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (N <= Q4K_GEMV_GEMM_MIN_N || N <= knobs().midn_max || !knobs().gemm_enabled) return Q4K_ERR_UNSUPPORTED;     // decode shapes quantize inside their launch
    if (M == 0) return Q4K_OK;
    if (!w_dev || !dst_dev || !workspace_dev) return Q4K_ERR_BAD_POINTER;
    if (workspace_bytes < q4k_mul_mat_workspace_bytes(K, N)) return Q4K_ERR_WORKSPACE;
    const int64_t nb = K / Q4K_QK_K;
    if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
    if (dst_col_stride_bytes < M * 4 || (dst_col_stride_bytes & 3) || !aligned16(workspace_dev)) return Q4K_ERR_BAD_SHAPE;
    const cudaError_t e = launch_gemm_q4k_f32(w_dev, M, K, w_row_stride_bytes, nullptr, N, 0, dst_dev, dst_col_stride_bytes,
                                              workspace_dev, (cudaStream_t)stream);
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

int q4k_mul_mat_peers_sync(const void *w_dev, int64_t M_local, int64_t K, int64_t w_row_stride_bytes,
                           const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                           float *const *dst_peers, int n_peers, int64_t dst_row0, int64_t dst_col_stride_bytes,
                           void *workspace_dev, size_t workspace_bytes,
                           void *const *flags_peers, int world, int rank, void *epoch_dev, void *done_dev, void *stream) {
    if (world < 1 || world > 8 || rank < 0 || rank >= world) return Q4K_ERR_BAD_SHAPE;
    if (!flags_peers || !epoch_dev || !done_dev) return Q4K_ERR_BAD_POINTER;
    PeerSyncReq req;
    for (int r = 0; r < 8; ++r) req.flags[r] = nullptr;
    for (int r = 0; r < world; ++r) {
        if (!flags_peers[r]) return Q4K_ERR_BAD_POINTER;
        req.flags[r] = static_cast<uint32_t *>(flags_peers[r]);
    }
    req.epoch = static_cast<uint32_t *>(epoch_dev); req.done = static_cast<uint32_t *>(done_dev);
    req.world = world; req.rank = rank; req.taken = false;
    t_sync_req = &req;
    const int rc = q4k_mul_mat_peers_ws(w_dev, M_local, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_peers, n_peers, dst_row0,
                                        dst_col_stride_bytes, workspace_dev, workspace_bytes, stream);
    t_sync_req = nullptr;
    if (rc != Q4K_OK) return rc;
    // no kernel carried the request (this rank owns no rows): the exchange runs as its own kernel
    if (!req.taken) return q4k_peer_signal_wait(flags_peers, world, rank, epoch_dev, stream);
    return Q4K_OK;
}

int q4k_mul_mat_peers(const void *w_dev, int64_t M_local, int64_t K, int64_t w_row_stride_bytes,
                      const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                      float *const *dst_peers, int n_peers, int64_t dst_row0, int64_t dst_col_stride_bytes, void *stream) {
    return q4k_mul_mat_peers_ws(w_dev, M_local, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_peers, n_peers, dst_row0,
                                dst_col_stride_bytes, nullptr, 0, stream);
}

int q4k_mul_mat_group(int n_mats, const void *const *w_dev, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                      const float *const *x_dev, int64_t N, int64_t x_row_stride_bytes, float *const *dst_dev, void *stream) {
    if (n_mats < 1 || n_mats > 4 || N < 1 || N > Q4K_GEMV_MAX_N) return Q4K_ERR_UNSUPPORTED;
    if (K <= 0 || K % Q4K_QK_K != 0 || !w_dev || !M || !x_dev || !dst_dev) return Q4K_ERR_BAD_SHAPE;
    const int64_t nb = K / Q4K_QK_K;
    if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15)) return Q4K_ERR_BAD_SHAPE;
    if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15)) return Q4K_ERR_BAD_SHAPE;
    for (int i = 0; i < n_mats; ++i) {
        if (M[i] < 0) return Q4K_ERR_BAD_SHAPE;
        if (!w_dev[i] || !x_dev[i] || !dst_dev[i] || !aligned16(w_dev[i]) || !aligned16(x_dev[i])) return Q4K_ERR_BAD_POINTER;
    }
    const cudaError_t e = launch_gemv_q4k_group(n_mats, w_dev, M, K, w_row_stride_bytes, x_dev, (int)N, x_row_stride_bytes, dst_dev,
                                                gemv_exact_for((int)N), knobs().pdl != 0, (cudaStream_t)stream);
    if (e == cudaErrorNotSupported) return Q4K_ERR_UNSUPPORTED;
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

int q4k_mul_mat_group_ws(int n_mats, const void *const *w_dev, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                         const float *x_dev, int64_t N, int64_t x_row_stride_bytes, float *const *dst_dev,
                         void *workspace_dev, size_t workspace_bytes, void *stream) {
    if (n_mats < 1 || n_mats > 4 || N < 1) return Q4K_ERR_UNSUPPORTED;
    if (K <= 0 || K % Q4K_QK_K != 0 || !w_dev || !M || !x_dev || !dst_dev) return Q4K_ERR_BAD_SHAPE;
    if (N <= Q4K_GEMV_MAX_N) {                      // decode: the group launch of the decode kernel (quantizer fused in)
        const float *xs[4] = {x_dev, x_dev, x_dev, x_dev};
        return q4k_mul_mat_group(n_mats, w_dev, M, K, w_row_stride_bytes, xs, N, x_row_stride_bytes, dst_dev, stream);
    }
    if (N <= knobs().midn_max || !knobs().gemm_enabled) return Q4K_ERR_UNSUPPORTED;      // the batched decode kernel takes one matrix at a time
    const int64_t nb = K / Q4K_QK_K;
    if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15)) return Q4K_ERR_BAD_SHAPE;
    if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
    if (!workspace_dev || !aligned16(workspace_dev)) return Q4K_ERR_BAD_POINTER;
    if (workspace_bytes < q4k_mul_mat_workspace_bytes(K, N)) return Q4K_ERR_WORKSPACE;
    int64_t strides[4];
    for (int i = 0; i < n_mats; ++i) {
        if (M[i] < 0) return Q4K_ERR_BAD_SHAPE;
        if (!w_dev[i] || !dst_dev[i] || !aligned16(w_dev[i])) return Q4K_ERR_BAD_POINTER;
        strides[i] = M[i] * 4;
    }
    const cudaError_t e = launch_gemm_q4k_f32_group(n_mats, w_dev, M, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_dev, strides,
                                                    workspace_dev, (cudaStream_t)stream);
    if (e == cudaErrorNotSupported) return Q4K_ERR_UNSUPPORTED;
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

int q4k_vec_dot_q4_K_q8_K(int n, float *s_dev, size_t bs, const void *vx_dev, size_t bx, const void *vy_dev, size_t by, int nrc) {
    (void)bs; (void)bx; (void)by;
    if (nrc != 1) return Q4K_ERR_UNSUPPORTED;      // the reference asserts nrc == 1 on x86 (arch/x86/quants.c:1743)
    if (n < 0 || n % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (!s_dev) return Q4K_ERR_BAD_POINTER;
    if (n == 0) {
        cudaError_t e0 = cudaMemset(s_dev, 0, 4);
        return e0 == cudaSuccess ? Q4K_OK : cuda_fail(e0);
    }
    if (!vx_dev || !vy_dev || !aligned16(vx_dev) || (reinterpret_cast<uintptr_t>(vy_dev) & 3u)) return Q4K_ERR_BAD_POINTER;
    const int64_t nb = n / Q4K_QK_K;
    cudaError_t e = launch_gemv_q4k_q8k(vx_dev, 1, n, nb * Q4K_BLOCK_Q4K, vy_dev, 1, nb * Q4K_BLOCK_Q8K, s_dev, 4, true, 0);
    if (e == cudaSuccess) e = cudaStreamSynchronize(0);
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

int q4k_mul_mat_dump(int kernel, const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                     const float *x_dev, int64_t N, int64_t x_row_stride_bytes, float *dst_dev,
                     int32_t *p_dev, int32_t *sumi_dev, int32_t *summ_dev,
                     void *workspace_dev, size_t workspace_bytes, void *stream) {
    if (M <= 0 || N <= 0 || K <= 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (kernel < Q4K_DUMP_DISPATCH || kernel > Q4K_DUMP_WIDE) return Q4K_ERR_UNSUPPORTED;
    if (!w_dev || !x_dev || !dst_dev || !sumi_dev || !summ_dev) return Q4K_ERR_BAD_POINTER;
    const int64_t nb = K / Q4K_QK_K;
    if (w_row_stride_bytes < nb * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15) || !aligned16(w_dev)) return Q4K_ERR_BAD_SHAPE;
    if (x_row_stride_bytes < K * 4 || (x_row_stride_bytes & 15) || !aligned16(x_dev)) return Q4K_ERR_BAD_SHAPE;
    if (workspace_bytes < q4k_mul_mat_workspace_bytes(K, N)) return Q4K_ERR_WORKSPACE;
    if (!workspace_dev || !aligned16(workspace_dev)) return Q4K_ERR_BAD_POINTER;
    cudaStream_t st = (cudaStream_t)stream;
    // which product kernel: the one q4k_mul_mat would launch for this shape, unless the caller names one
    const bool wide = kernel == Q4K_DUMP_WIDE || (kernel == Q4K_DUMP_DISPATCH && N >= knobs().wide_min && N <= knobs().midn_max &&
                                                  !(N <= Q4K_GEMV_MAX_N && gemv_exact_for((int)N)) && !(N == 2 && K < 4096));
    if (wide) {
        if (!p_dev || !aligned16(workspace_dev)) return Q4K_ERR_BAD_POINTER;
        const IntDump d{p_dev, sumi_dev, summ_dev};
        const cudaError_t e = launch_gemv_q4k_wide(w_dev, M, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_dev, M * 4, workspace_dev,
                                                   knobs().pdl != 0, st, &d);
        if (e == cudaErrorNotSupported) return Q4K_ERR_UNSUPPORTED;
        return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
    }
    const bool gemm = kernel == Q4K_DUMP_GEMM || (kernel == Q4K_DUMP_DISPATCH && N > Q4K_GEMV_GEMM_MIN_N && knobs().gemm_enabled);
    if (gemm) {
        if (N <= Q4K_GEMV_MAX_N) return Q4K_ERR_UNSUPPORTED;
        const IntDump d{nullptr, sumi_dev, summ_dev};
        const cudaError_t e = launch_gemm_q4k_f32(w_dev, M, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_dev, M * 4,
                                                  workspace_dev, st, &d);
        return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
    }
    if (!p_dev) return Q4K_ERR_BAD_POINTER;
    const bool exact = kernel == Q4K_DUMP_BULK_EXACT || (kernel == Q4K_DUMP_DISPATCH && gemv_exact_for((int)N));
    const int which = kernel == Q4K_DUMP_DISPATCH ? 0 : (kernel == Q4K_DUMP_TENSOR_PIPE ? 1 : (kernel == Q4K_DUMP_BIG ? 3 : 2));
    int64_t step = Q4K_GEMV_MAX_N;
    for (int64_t n0 = 0; n0 < N;) {
        const int nn = (int)((N - n0) < step ? (N - n0) : step);
        const IntDump d{p_dev + n0 * M * nb * 8, sumi_dev + n0 * M * nb, summ_dev + n0 * M * nb};
        const float *xs = (const float *)((const char *)x_dev + n0 * x_row_stride_bytes);
        const void *xq = nullptr;
        int64_t xstride = x_row_stride_bytes;
        if ((int64_t)nn * K > knobs().fuse_limit) {             // as q4k_mul_mat: long activations go through the standalone quantizer
            const cudaError_t eq = launch_quantize_q8_K(xs, workspace_dev, K, nn, x_row_stride_bytes, st);
            if (eq != cudaSuccess) return cuda_fail(eq);
            xq = workspace_dev; xs = nullptr; xstride = nb * Q4K_BLOCK_Q8K;
        }
        const cudaError_t e = launch_gemv_q4k_q8k_tma(w_dev, M, K, w_row_stride_bytes, xq, xs, nn, xstride, dst_dev + n0 * M, M * 4,
                                                      exact, 0, knobs().pdl != 0, st, nullptr, 0, 0, &d, which);
        if (e == cudaErrorNotSupported && nn > 1) { step = nn / 2; continue; }     // activations too long for one residency: fewer columns
        if (e == cudaErrorNotSupported) return Q4K_ERR_UNSUPPORTED;
        if (e != cudaSuccess) return cuda_fail(e);
        n0 += nn;
    }
    return Q4K_OK;
}

int q4k_mul_mat_host_async(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                           const float *x_host, int64_t N, float *dst_host,
                           void *x_stage_dev, float *dst_stage_dev,
                           void *workspace_dev, size_t workspace_bytes, void *stream) {
    if (M < 0 || N < 0 || K < 0) return Q4K_ERR_BAD_SHAPE;
    if (M == 0 || N == 0) return Q4K_OK;
    if (!x_host || !dst_host || !x_stage_dev || !dst_stage_dev) return Q4K_ERR_BAD_POINTER;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaSuccess;
    if (K > 0) e = cudaMemcpyAsync(x_stage_dev, x_host, (size_t)N * K * 4, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return cuda_fail(e);
    const int rc = q4k_mul_mat(w_dev, M, K, w_row_stride_bytes, (const float *)x_stage_dev, N, K * 4,
                               dst_stage_dev, M * 4, workspace_dev, workspace_bytes, stream);
    if (rc != Q4K_OK) return rc;
    e = cudaMemcpyAsync(dst_host, dst_stage_dev, (size_t)N * M * 4, cudaMemcpyDeviceToHost, st);
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

int q4k_mul_mat_host(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                     const float *x_host, int64_t N, float *dst_host,
                     void *x_stage_dev, float *dst_stage_dev,
                     void *workspace_dev, size_t workspace_bytes, void *stream) {
    const int rc = q4k_mul_mat_host_async(w_dev, M, K, w_row_stride_bytes, x_host, N, dst_host, x_stage_dev, dst_stage_dev,
                                          workspace_dev, workspace_bytes, stream);
    if (rc != Q4K_OK || M <= 0 || N <= 0) return rc;
    const cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
    return e == cudaSuccess ? Q4K_OK : cuda_fail(e);
}

}  // extern "C"
