// noalloc.cu — q4k_mul_mat_noalloc: the decode mul_mat (N <= 8) for callers that cannot allocate.
//
// q4k_mul_mat needs a caller-allocated workspace for the Q8_K copy of the activations (the reference's params->wdata): the
// standalone quantizer kernel writes it, the decode kernel reads it (activations above the fuse limit, e.g. K = 28672).  Code
// under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc have nowhere to get that workspace from.
// This entry point brings its own: a file-scope static device array of Q4K_NOALLOC_SCRATCH_BYTES, resident on every GPU that
// loads the library until the process exits.  It calls no cudaMalloc / cudaFree / allocating library routine.
//
// No counterpart in the reference.
//
// The array's NAME is a host-side stand-in in host code, not a device address: the device address is fetched once per device
// (with that device current) by cudaGetSymbolAddress and that pointer is what the kernels get.
// Every call on a device shares the one array, which keeps the previous call's contents.  Nothing in it has to start at
// zero: the quantizer overwrites every Q8_K block the decode kernel then reads.  Calls that overlap on the GPU (two streams,
// two threads on one device) race on it — serialise them, or use q4k_mul_mat with a workspace of your own.
#include "../../include/q4k_b200.h"

#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

namespace q4k {
int note_cuda_error(cudaError_t e);      // c_api.cu
}

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[Q4K_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
void *g_scratch_ptr[MAX_DEVICES] = {nullptr};
std::mutex g_scratch_mu;

// device address of the scratch array on the CURRENT device (looked up once per device)
int scratch_for_current_device(void **out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return q4k::note_cuda_error(e);
    if (dev < 0 || dev >= MAX_DEVICES) return Q4K_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return q4k::note_cuda_error(e);
        if (!p) return Q4K_ERR_BAD_POINTER;
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return Q4K_OK;
}

}  // namespace

extern "C" int q4k_mul_mat_noalloc(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                   const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                                   float *dst_dev, int64_t dst_col_stride_bytes, void *stream) {
    if (M < 0 || N < 0 || K < 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (N > Q4K_GEMV_MAX_N) return Q4K_ERR_STATIC_LIMIT;                       // decode shapes only
    const size_t need = q4k_mul_mat_workspace_bytes(K, N);                     // N * K/256 * 292 for N = 1; the eight-column staged copy of gemv_q4k_wide.cu for N = 2..8
    if (need > (size_t)Q4K_NOALLOC_SCRATCH_BYTES) return Q4K_ERR_STATIC_LIMIT;
    if (M == 0 || N == 0) return Q4K_OK;
    void *scratch = nullptr;
    const int rc = scratch_for_current_device(&scratch);
    if (rc != Q4K_OK) return rc;
    return q4k_mul_mat(w_dev, M, K, w_row_stride_bytes, x_dev, N, x_row_stride_bytes, dst_dev, dst_col_stride_bytes,
                       scratch, (size_t)Q4K_NOALLOC_SCRATCH_BYTES, stream);
}
