// q4k_kernels.h — internal launcher interface between the kernel TUs and the C ABI (c_api.cu).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "knobs.h"

namespace q4k {

void count_launch();          // bumps the process-wide kernel launch counter (q4k_kernel_launches)

// parity tooling: where a DUMP instantiation of a product kernel writes the integers it computed
// (p[((n*M + m)*nb + b)*8 + j], sumi / summ[(n*M + m)*nb + b]; p may be null for the GEMM, which never forms p[j])
struct IntDump { int32_t *p, *sumi, *summ; };

// LL exchange of the row-sharded multi-GPU path (q4k_mul_mat_peers_ll): every result leaves the kernel as an 8-byte
// {value, epoch} pair, stored (one atomic 8-byte NVLink store, no fence) into the LL buffer of EVERY rank; the tail of the same
// launch polls this rank's own LL buffer until every rank's pairs of THIS call have arrived and writes them out as plain floats.
// One launch = compute + all-gather + completion: no flag kernel, no system-scope fence.
// Back-pressure: a rank may overwrite half h with call k + 2 only after EVERY rank has finished reading call k out of it.  Data
// pairs give that for ranks that own rows (a rank's call-(k+1) pairs leave after its call k has completed, and everybody's
// call k + 1 waits for them); a rank that owns NO rows sends nothing, so every rank also sends one "presence" pair per call
// ({0, epoch} into slot [half][rank] behind the data) and every tail waits for all of them.
// LL buffer of a rank: [2 halves][n_max columns][M_full rows] pairs + [2 halves][8] presence pairs; the call's epoch = cnt[0] + 1 (device-side, so a CUDA
// graph can replay the launch), the half it uses = epoch & 1; cnt[1] = CTA arrival counter (the last CTA publishes cnt[0]).
struct LLExchange {
    void *peers[8];          // every rank's LL buffer as mapped on THIS GPU (own one included), indexed by rank
    void *local;             // this rank's own LL buffer
    int64_t half_pairs;      // pairs per half = n_max * M_full
    int64_t M_full;          // rows of the full matrix
    uint32_t *cnt;           // two zero-initialised words in this rank's device memory
    float *dst;              // plain F32 result [N][dst_stride] in this rank's memory, complete when the launch completes
    int64_t dst_stride;      // floats between columns
    int world, rank;
};

cudaError_t launch_quantize_q8_K(const float *x, void *y, int64_t k, int64_t nrows, int64_t x_row_stride_bytes,
                                 cudaStream_t stream);

// GEMV, 1 <= N <= 8.  exact_order: reproduce the reference's AVX2 float accumulation order bit for bit.
cudaError_t launch_gemv_q4k_q8k(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                const void *xq, int N, int64_t xq_row_stride_bytes,
                                float *dst, int64_t dst_col_stride_bytes, bool exact_order, cudaStream_t stream);

// the TMA-bulk staged implementation (gemv_q4k_tma.cu).  Activations: xq (Q8_K blocks) or, when xq == nullptr,
// xf (F32 rows, quantized inside the kernel).  stages <= 0 = choose; pdl = programmatic dependent launch.
// cudaErrorNotSupported if the Q8_K activations do not fit in shared memory.
// peer_dst / n_peers / dst_row0: row-sharded multi-GPU — the result rows are stored at row offset dst_row0 into the
// output buffer of every listed GPU (peer-mapped pointers; peer_dst[0] should be the local one).
cudaError_t launch_gemv_q4k_q8k_tma(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                    const void *xq, const float *xf, int N, int64_t x_row_stride_bytes,
                                    float *dst, int64_t dst_col_stride_bytes, bool exact_order, int stages,
                                    bool pdl, cudaStream_t stream,
                                    float *const *peer_dst = nullptr, int n_peers = 0, int64_t dst_row0 = 0,
                                    const IntDump *dump = nullptr, int dump_kernel = 0, const LLExchange *ll = nullptr);

// up to 4 matrices (same K, N) in one launch; see gemv_q4k_tma.cu
cudaError_t launch_gemv_q4k_group(int nmat, const void *const *w, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                                  const float *const *x, int N, int64_t x_row_stride_bytes, float *const *dst,
                                  bool exact_order, bool pdl, cudaStream_t stream);

// the tensor-pipe decode kernel (gemv_q4k_mma.cu): tolerance float order, N <= 8; see there for the arguments
cudaError_t launch_gemv_q4k_mma(int nmat, const void *const *w, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                                const void *xq, const float *const *xf, int N, int64_t x_row_stride_bytes,
                                float *const *dst, int64_t dst_col_stride_bytes, float *const *peer_dst, int n_peers,
                                int64_t dst_row0, bool pdl, cudaStream_t stream, const IntDump *dump = nullptr,
                                const LLExchange *ll = nullptr);

// the tail of an LL exchange as a kernel of its own: for a rank that owns no rows of a matrix (its launch has nothing to compute)
cudaError_t launch_ll_gather(const LLExchange &ll, int N, bool pdl, cudaStream_t stream);

// GEMM on tcgen05 (N > 8).  Returns cudaErrorNotSupported when the shape cannot take the tensor path.
cudaError_t launch_gemm_q4k_q8k(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                const void *xq, int64_t N, int64_t xq_row_stride_bytes,
                                float *dst, int64_t dst_col_stride_bytes, cudaStream_t stream);

// F32 activations, N > 8: planar Q8 quantizer + tcgen05 GEMM; `workspace` >= gemm_workspace_bytes(K, N)
size_t gemm_workspace_bytes(int64_t K, int64_t N);
cudaError_t launch_gemm_q4k_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                const float *x, int64_t N, int64_t x_row_stride_bytes,
                                float *dst, int64_t dst_col_stride_bytes, void *workspace, cudaStream_t stream,
                                const IntDump *dump = nullptr);

// batched decode on the tensor pipe (gemv_q4k_wide.cu): F32 activations, 8 < N; quantizes into `workspace`
// (wide_workspace_bytes) in its own staged layout and streams weights once per group of up to WIDE_MAX_COLS columns
constexpr int WIDE_MAX_COLS = 32;
size_t wide_workspace_bytes(int64_t K, int64_t N);
cudaError_t launch_gemv_q4k_wide(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x, int64_t N,
                                 int64_t x_row_stride_bytes, float *dst, int64_t dst_col_stride_bytes, void *workspace, bool pdl,
                                 cudaStream_t stream, const IntDump *dump = nullptr);

// streamed single-column decode (gemv_q4k_n1s.cu): F32 activation row, quantized into `workspace` (n1s_workspace_bytes) and
// streamed next to the weights; same bits as gemv_q4k_mma.cu at N == 1
size_t n1s_workspace_bytes(int64_t K);
// ll != nullptr: the row-sharded form — rows leave as LL pairs at row offset dst_row0 and the launch's tail gathers (dst unused)
cudaError_t launch_gemv_q4k_n1s(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x, float *dst,
                                void *workspace, bool pdl, cudaStream_t stream, const LLExchange *ll = nullptr, int64_t dst_row0 = 0);

// up to four matrices with the same K on ONE activation tensor as one GEMM launch (one quantizer pass, one grid over all row tiles)
cudaError_t launch_gemm_q4k_f32_group(int nmat, const void *const *w, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                                      const float *x, int64_t N, int64_t x_row_stride_bytes, float *const *dst,
                                      const int64_t *dst_col_stride_bytes, void *workspace, cudaStream_t stream,
                                      const IntDump *dump = nullptr);

// Q6_K decode GEMV (gemv_q6k.cu), 1 <= N <= 8; dump != nullptr also writes the per-superblock integer sums
cudaError_t launch_gemv_q6k(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                            const void *xq, const float *xf, int N, int64_t x_row_stride_bytes,
                            float *dst, int64_t dst_col_stride_bytes, int32_t *dump, bool pdl, cudaStream_t stream);

// Q6_K prefill GEMM on tcgen05 kind::f16 (gemm_q6k_tc.cu): F32 activations, quantized into `workspace` (gemm_q6k_workspace_bytes)
// as fp16 tiles; dump_sumi != nullptr runs the parity instantiation that also writes every superblock's integer sum
size_t gemm_q6k_workspace_bytes(int64_t K, int64_t N);
cudaError_t launch_gemm_q6k_f32(const void *w, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x, int64_t N,
                                int64_t x_row_stride_bytes, float *dst, int64_t dst_col_stride_bytes, void *workspace,
                                cudaStream_t stream, int32_t *dump_sumi = nullptr);

// request handed from the C ABI (q4k_mul_mat_peers_sync) to whichever launcher ends up launching the single decode kernel
// of the call: the kernel's last CTA then performs the cross-GPU flag exchange itself (gemv_stage.cuh peer_sync_tail)
struct PeerSyncReq {
    uint32_t *flags[8];     // every rank's flag row (peer-mapped), indexed by rank
    uint32_t *epoch, *done; // this rank's call counter and CTA arrival counter (device memory, zero-initialised)
    int world, rank;
    bool taken;             // set by the launcher once a kernel carrying the request has been launched
};
extern thread_local PeerSyncReq *t_sync_req;

void add_launches(int64_t n);  // launches replayed through a CUDA graph
int64_t launch_count();
int sm_count();               // multiprocessors of the current device (cached)

}  // namespace q4k
