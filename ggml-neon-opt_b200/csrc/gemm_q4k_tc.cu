// gemm_q4k_tc.cu — prefill GEMM (N > 8) on tcgen05 kind::i8 with TMEM accumulators.
// Placeholder until the tensor-core kernel lands: reports "not supported" so the C ABI takes the
// GEMV-per-8-columns route (same GPU arithmetic, not a CPU fallback).
#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {

cudaError_t launch_gemm_q4k_q8k(const void *, int64_t, int64_t, int64_t, const void *, int64_t, int64_t,
                                float *, int64_t, cudaStream_t) {
    return cudaErrorNotSupported;
}

}  // namespace q4k
