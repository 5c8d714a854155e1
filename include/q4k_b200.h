/*
 * q4k_b200.h — C ABI of the B200-native Q4_K x Q8_K mul_mat path (libq4k_b200.so).
 *
 * This is the drop-in boundary under the ggml backend plugin (ggml-neon-opt_b200/backend/) and the
 * thing tests/bench call through ctypes.  Plain pointers and sizes only; every pointer named *_dev is a
 * CUDA device pointer, `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 * Every entry returns 0 on success or a negative Q4K_ERR_* code; nothing throws; launches are
 * asynchronous on `stream` unless the comment says "synchronous".
 *
 * Byte layouts are the reference's, unchanged ("L/" = external/llama.cpp/ in the reference):
 *   block_q4_K  144 B / 256 weights      L/ggml/src/ggml-common.h:291-306
 *   block_q8_K  292 B / 256 activations  L/ggml/src/ggml-common.h:338-344
 * Matrix convention is ggml's (L/ggml/src/ggml-cpu/ggml-cpu.c:1221-1224): weights src0 = M rows of K,
 * activations src1 = N rows of K floats, dst = N columns of M floats (M fastest).
 */
#ifndef Q4K_B200_H
#define Q4K_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define Q4K_QK_K          256
#define Q4K_BLOCK_Q4K     144
#define Q4K_BLOCK_Q8K     292
#define Q4K_BLOCK_Q6K     210    /* block_q6_K, L/ggml/src/ggml-common.h:326-336 */
#define Q4K_GEMV_MAX_N    8      /* batch sizes served by the HBM-bound GEMV; larger N takes the GEMM */

enum {
    Q4K_OK               =  0,
    Q4K_ERR_BAD_SHAPE    = -1,   /* K % 256 != 0, negative sizes, strides too small / misaligned */
    Q4K_ERR_BAD_POINTER  = -2,
    Q4K_ERR_WORKSPACE    = -3,   /* workspace smaller than q4k_mul_mat_workspace_bytes() */
    Q4K_ERR_CUDA         = -4,   /* a CUDA runtime call failed; q4k_last_cuda_error() has the code */
    Q4K_ERR_UNSUPPORTED  = -5,   /* nrc not in {1}, device is not sm_100 ... */
    Q4K_ERR_STATIC_LIMIT = -6,   /* q4k_mul_mat_noalloc: the shape needs more scratch than the library's static array holds */
};

/* library / device ------------------------------------------------------------------------------- */
int         q4k_abi_version(void);                 /* bumps when a signature changes */
const char *q4k_build_info(void);                  /* "sm_100a nvcc 12.9 ..." */
int         q4k_device_ok(int device);             /* 1 if `device` is compute capability 10.x */
int         q4k_last_cuda_error(void);             /* last cudaError_t seen by this library, per thread */
int64_t     q4k_kernel_launches(void);             /* kernels launched by this library so far (process-wide) */

/* device memory / streams for host code above this ABI that is compiled without CUDA headers (the ggml backend
 * plugin: buffer_type_i.alloc_buffer, buffer_i.set_tensor / get_tensor / memset_tensor / clear / cpy_tensor and
 * backend_i.synchronize of L/ggml/src/ggml-backend-impl.h:17-126).  All act on `device`; copies with a NULL
 * stream are synchronous, with a stream they are queued on it (host memory should be pinned then). */
int     q4k_device_count(void);                                        /* number of CUDA devices (plain ordinals; check each with q4k_device_ok) */
int     q4k_device_info(int device, char *name, size_t name_len, size_t *free_bytes, size_t *total_bytes, int *sm_count);
void   *q4k_mem_alloc(int device, size_t bytes);                       /* NULL on failure (q4k_last_cuda_error) */
int     q4k_mem_free(int device, void *ptr_dev);
int     q4k_mem_set(int device, void *dst_dev, int value, size_t bytes, void *stream);
int     q4k_mem_h2d(int device, void *dst_dev, const void *src_host, size_t bytes, void *stream);
int     q4k_mem_d2h(int device, void *dst_host, const void *src_dev, size_t bytes, void *stream);
int     q4k_mem_d2d(int device, void *dst_dev, const void *src_dev, size_t bytes, void *stream);
void   *q4k_host_alloc_pinned(size_t bytes);                           /* page-locked host memory for async copies */
int     q4k_host_free_pinned(void *ptr_host);
void   *q4k_stream_create(int device);                                 /* cudaStream_t as void*; NULL on failure */
int     q4k_stream_destroy(int device, void *stream);
int     q4k_stream_synchronize(int device, void *stream);
int     q4k_set_device(int device);                                    /* makes `device` current for the calling thread */

/* Captured launch sequences: what backend_i.graph_plan_create / graph_plan_compute (L/ggml/src/ggml-backend-impl.h:100-108)
 * and the plugin's cached graph_compute are made of.  Between begin and end every q4k_* launch on `stream` is recorded (CUDA
 * stream capture, programmatic-dependent-launch edges included) instead of executed; end instantiates the graph (keep = 0:
 * abandon it), launch replays it on a stream.  Nothing may synchronize or allocate between begin and end. */
typedef struct q4k_graph q4k_graph;
int  q4k_graph_begin_capture(void *stream);
int  q4k_graph_end_capture(void *stream, int keep, q4k_graph **graph_out);
int  q4k_graph_launch(q4k_graph *graph, void *stream);
void q4k_graph_destroy(q4k_graph *graph);

/* replaces quantize_row_q8_K (L/ggml/src/ggml-quants.c:2555-2592, bound as from_float of Q8_K at
 * L/ggml/src/ggml-cpu/ggml-cpu.c:355-357) for `nrows` rows of k floats; row r is read at
 * x_dev + r*x_row_stride_bytes and written to y_dev + r*(k/256*292).  Bit-exact, except that bsums of
 * all-zero blocks are written as 0 (the reference leaves them unwritten). */
int q4k_quantize_q8_K(const float *x_dev, void *y_dev, int64_t k, int64_t nrows,
                      int64_t x_row_stride_bytes, void *stream);

/* replaces ggml_compute_forward_mul_mat for Q4_K src0 / F32 src1 (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401).
 * workspace_dev holds the Q8_K copy of src1 (the reference's params->wdata, :1267-1307). */
size_t q4k_mul_mat_workspace_bytes(int64_t K, int64_t N);
int q4k_mul_mat(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                float *dst_dev, int64_t dst_col_stride_bytes,
                void *workspace_dev, size_t workspace_bytes, void *stream);

/* q4k_mul_mat for callers that CANNOT allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc): the same decode mul_mat (N <= 8) with the Q8_K scratch in a static device array inside the library
 * instead of a caller-supplied workspace.  No counterpart in the reference.
 *   - calls no cudaMalloc / cudaFree / allocating library routine; inputs and outputs are still the caller's;
 *   - limit: N <= 8 and q4k_mul_mat_workspace_bytes(K, N) <= Q4K_NOALLOC_SCRATCH_BYTES (1 MiB: K up to 112640 for N = 2..8 (N = 2: from K = 4096), whose
 *     staged Q8_K copy is padded to eight columns, 9.25 B per K; 292 B per 256 activations for N = 1), else Q4K_ERR_STATIC_LIMIT; every GPU that loads the library holds the 1 MiB until the process exits;
 *   - runs on the CURRENT device; results are bit-identical to q4k_mul_mat on the same inputs;
 *   - all calls on a device share the one array, which keeps the last call's contents (nothing in it needs to start at zero:
 *     the quantizer overwrites every block the decode kernel reads).  Calls that OVERLAP on the GPU (two streams, two threads
 *     on one device) race on it: serialise them, or use q4k_mul_mat with your own workspace. */
#define Q4K_NOALLOC_SCRATCH_BYTES (1 << 20)
int q4k_mul_mat_noalloc(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                        const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                        float *dst_dev, int64_t dst_col_stride_bytes, void *stream);

/* prefill (N > 32, "midn_max"): the mul_mat in front of this one (same stream, same workspace) read the SAME activations — attn_q /
 * attn_k / attn_v or ffn_gate / ffn_up of a layer, separate ggml_mul_mat nodes on one src1 in the reference's graph — so their
 * quantized copy is still in the workspace: run the GEMM only.  The reference quantizes src1 once per node
 * (L/ggml/src/ggml-cpu/ggml-cpu.c:1267-1307).  Q4K_ERR_UNSUPPORTED for N <= 32 (call q4k_mul_mat). */
int q4k_mul_mat_reuse(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes, int64_t N,
                      float *dst_dev, int64_t dst_col_stride_bytes, void *workspace_dev, size_t workspace_bytes, void *stream);

/* row-sharded multi-GPU decode (N <= 8): this GPU owns weight rows [dst_row0, dst_row0 + M_local) of a larger matrix.
 * Runs quantize + GEMV in ONE launch and stores every result at row dst_row0 + m of column n into the output buffer
 * of EVERY GPU in dst_peers[0..n_peers) (device pointers valid on this GPU: its own buffer and NVLink peer-mapped
 * ones, e.g. CUDA IPC / symmetric memory) — the all-gather of SURVEY.md §8e fused into the GEMV epilogue.  The
 * caller orders the peers' reads after all ranks' launches (a cross-GPU barrier).  What it replaces in the
 * reference: the dst-slice peer copies of ggml-cuda's row split (L/ggml/src/ggml-cuda/ggml-cuda.cu:1720). */
int q4k_mul_mat_peers(const void *w_dev, int64_t M_local, int64_t K, int64_t w_row_stride_bytes,
                      const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                      float *const *dst_peers, int n_peers, int64_t dst_row0, int64_t dst_col_stride_bytes, void *stream);

/* q4k_mul_mat_peers with a Q8_K workspace (>= q4k_mul_mat_workspace_bytes(K, N)): long activation rows (N*K above the
 * fuse limit) are quantized once by a kernel of their own instead of once per CTA, as q4k_mul_mat does. */
int q4k_mul_mat_peers_ws(const void *w_dev, int64_t M_local, int64_t K, int64_t w_row_stride_bytes,
                         const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                         float *const *dst_peers, int n_peers, int64_t dst_row0, int64_t dst_col_stride_bytes,
                         void *workspace_dev, size_t workspace_bytes, void *stream);

/* Row-sharded multi-GPU decode with the LL exchange — compute, all-gather and completion in ONE launch, without a flag kernel
 * and without a system-scope fence.  Every result leaves the kernel as an 8-byte {value, epoch} pair stored (one atomic NVLink
 * store) into the LL buffer of EVERY rank; the tail of the same launch polls this rank's own LL buffer until the pairs of this
 * call have arrived from everybody and writes them to dst_dev as plain floats: when the launch completes, dst_dev (this
 * GPU's memory, N columns of M_full rows at dst_col_stride_bytes) holds the full result.  What it replaces in the reference:
 * the event + cudaMemcpyPeerAsync gather of a row split (L/ggml/src/ggml-cuda/ggml-cuda.cu:1720-1760).
 *   ll_peers[r]  rank r's LL buffer as mapped on this GPU (q4k_ll_buffer_bytes(M_full, n_max) bytes, zero-initialised once,
 *                the same allocation on every rank: [2 halves][n_max][M_full] data pairs + [2][8] presence pairs);
 *   cnt_dev      two zero-initialised uint32 in this rank's memory (call counter, CTA arrivals).  The epoch of a call and the
 *                half it uses are derived from the counter ON THE DEVICE, so the launch can be captured into a CUDA graph and
 *                replayed; every rank must make the same sequence of calls on its copy of the object.
 * A rank with M_local == 0 still calls (it only gathers).  N <= 8. */
size_t q4k_ll_buffer_bytes(int64_t M_full, int64_t n_max);
int q4k_mul_mat_peers_ll(const void *w_dev, int64_t M_local, int64_t K, int64_t w_row_stride_bytes,
                         const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                         void *const *ll_peers, int world, int rank, int64_t M_full, int64_t n_max, int64_t dst_row0,
                         float *dst_dev, int64_t dst_col_stride_bytes, void *cnt_dev,
                         void *workspace_dev, size_t workspace_bytes, void *stream);

/* q4k_mul_mat_peers_ws + the cross-GPU step in ONE kernel: the decode kernel's last CTA publishes this rank's epoch into
 * every peer's flag row and waits for theirs, so when the launch completes every rank's rows have landed in this rank's
 * buffer — compute, all-gather and barrier in one launch.  Measured on B200 it is SLOWER than q4k_mul_mat_peers_ws followed by
 * q4k_peer_signal_wait (2 GPUs, 14336x8192 per rank: 25.3 vs 20.9 us per call): kept as an option, not the default.  flags_peers
 * / world / rank / epoch_dev as for q4k_peer_signal_wait; done_dev = one more zero-initialised uint32 in this rank's
 * device memory (CTA arrival counter).  Falls back to a separate q4k_peer_signal_wait launch when this rank owns no rows. */
int q4k_mul_mat_peers_sync(const void *w_dev, int64_t M_local, int64_t K, int64_t w_row_stride_bytes,
                           const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                           float *const *dst_peers, int n_peers, int64_t dst_row0, int64_t dst_col_stride_bytes,
                           void *workspace_dev, size_t workspace_bytes,
                           void *const *flags_peers, int world, int rank, void *epoch_dev, void *done_dev, void *stream);

/* the cross-GPU step after q4k_mul_mat_peers: "all ranks' result stores have landed everywhere".  flags_peers[r] = the
 * peer-mapped pointer to rank r's flag buffer (world x uint32, zero-initialised once, same symmetric allocation on every
 * rank), epoch_dev = a zero-initialised uint32 in this rank's device memory.  One small kernel: bumps the epoch, publishes
 * it into every peer's flags (system-scope release), waits for every peer's epoch (system-scope acquire).  Every rank must
 * call it the same number of times.  Graph-capturable; no host synchronisation.  Replaces the ncclAllGather / event +
 * cudaMemcpyPeerAsync gather of a row split (L/ggml/src/ggml-cuda/ggml-cuda.cu:1720-1760). */
int q4k_peer_signal_wait(void *const *flags_peers, int world, int rank, void *epoch_dev, void *stream);

/* horizontally fused decode launch: n_mats (<= 4) independent mul_mats with the same K and N (<= 8) in ONE kernel
 * launch — e.g. attn_q / attn_k / attn_v or ffn_gate / ffn_up of one layer, which the reference runs as separate
 * ggml_compute_forward_mul_mat nodes.  x_dev[i] may repeat: matrices reading the same activations share one quantized
 * copy.  Results are dense: dst_dev[i] is [N][M[i]].  Q4K_ERR_UNSUPPORTED when the shapes do not fit one launch. */
int q4k_mul_mat_group(int n_mats, const void *const *w_dev, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                      const float *const *x_dev, int64_t N, int64_t x_row_stride_bytes, float *const *dst_dev, void *stream);
/* The same for ANY batch size when every matrix reads the SAME activation tensor x_dev: N <= 8 -> q4k_mul_mat_group; N above the
 * batched decode kernel's range (32) -> one quantizer pass + ONE tensor-core GEMM launch whose grid covers the row tiles of all
 * matrices (a prefill launch costs ~20 us whatever its size: TinyLlama's attn_k / attn_v are 8 CTAs each when launched alone);
 * in between Q4K_ERR_UNSUPPORTED (run them one by one).  dst_dev[i] = dense [N][M_i].  Results equal the one-by-one results bit
 * for bit unless the grouped grid changes the GEMM's split-K choice (then within 1e-5). */
int q4k_mul_mat_group_ws(int n_mats, const void *const *w_dev, const int64_t *M, int64_t K, int64_t w_row_stride_bytes,
                         const float *x_dev, int64_t N, int64_t x_row_stride_bytes, float *const *dst_dev,
                         void *workspace_dev, size_t workspace_bytes, void *stream);

/* Graph plans — the backend_i.graph_plan_create / graph_plan_compute pair of L/ggml/src/ggml-backend-impl.h:100-108 for
 * this path: a fixed sequence of Q4_K mul_mats (one decode token's worth, say) recorded once, replayed as ONE CUDA
 * graph with programmatic-dependent-launch edges.  Inputs (F32 activations, one buffer per input_id) and outputs
 * (dense [N][M] per mul_mat) are packed into two contiguous device buffers owned by the plan, so a host caller moves
 * one block in and one block out per replay.  Consecutive mul_mats with the same input_id, K and N <= 8 are fused
 * into one launch (q4k_mul_mat_group).  All functions return 0 / a negative Q4K_ERR_* code unless noted. */
typedef struct q4k_plan q4k_plan;
q4k_plan *q4k_plan_create(int device);                                   /* NULL on failure */
int       q4k_plan_add_mul_mat(q4k_plan *plan, const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                               int64_t N, int input_id);                 /* returns the mul_mat's index (>= 0) */
int       q4k_plan_add_mul_mat_q6k(q4k_plan *plan, const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                                   int64_t N, int input_id);             /* same, block_q6_K weights (q6k_mul_mat) */
int       q4k_plan_finalize(q4k_plan *plan);                             /* allocates staging, captures the graph */
size_t    q4k_plan_input_bytes(const q4k_plan *plan);
size_t    q4k_plan_output_bytes(const q4k_plan *plan);
int64_t   q4k_plan_input_offset(const q4k_plan *plan, int input_id);     /* byte offset of [N][K] floats; -1 if unknown */
int64_t   q4k_plan_output_offset(const q4k_plan *plan, int index);       /* byte offset of [N][M] floats */
void     *q4k_plan_input_dev(q4k_plan *plan);
void     *q4k_plan_output_dev(q4k_plan *plan);
int64_t   q4k_plan_launches(const q4k_plan *plan);                       /* kernel launches per replay */
int       q4k_plan_run(q4k_plan *plan, void *stream);                    /* inputs already in the plan's device buffer */
int       q4k_plan_run_host(q4k_plan *plan, const void *in_host, void *out_host, void *stream);  /* H2D, replay, D2H; async */
void      q4k_plan_destroy(q4k_plan *plan);

/* same, src1 already Q8_K (vec_dot_type == src1->type branch, ggml-cpu.c:1267): N rows of K/256 blocks */
int q4k_mul_mat_q8(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                   const void *xq_dev, int64_t N, int64_t xq_row_stride_bytes,
                   float *dst_dev, int64_t dst_col_stride_bytes, void *stream);

/* ggml_vec_dot_t signature (L/ggml/include/ggml-cpu.h:109; ggml_vec_dot_q4_K_q8_K at
 * L/ggml/src/ggml-cpu/arch/x86/quants.c:1741) over DEVICE pointers: n = K, s = result, nrc must be 1
 * (bs/bx/by unused as in the reference's nrc==1 path).  Float accumulation follows the reference's
 * AVX2 lane order exactly, so the result is bit-identical to the x86 build.  Synchronous. */
int q4k_vec_dot_q4_K_q8_K(int n, float *s_dev, size_t bs, const void *vx_dev, size_t bx,
                          const void *vy_dev, size_t by, int nrc);

/* debug / parity: q4k_mul_mat through a DUMP instantiation of the PRODUCT kernel itself — the same
 * code, compiled with a flag that also writes out the integers it computed (the generic reference materialises them:
 * L/ggml/src/ggml-cpu/quants.c:575-611, aux32 -> sumi, and the mins sum):
 *   p_dev   [((n*M + m)*nb + b)*8 + j]  int32  dot product of sub-block j of superblock b (row m, column n)
 *   sumi_dev[(n*M + m)*nb + b]          int32  sum_j sc[j] * p[j]
 *   summ_dev[(n*M + m)*nb + b]          int32  sum_j m[j] * bsum32[j]
 * dst_dev = dense [N][M] fp32 results of the same launch.  `kernel` picks the product kernel: */
enum {
    Q4K_DUMP_DISPATCH    = 0,    /* whatever q4k_mul_mat launches for this shape (decode kernels, the batched decode kernel for 2..32 columns, the GEMM above) */
    Q4K_DUMP_TENSOR_PIPE = 1,    /* gemv_q4k_mma_kernel (the default decode kernel; p[j] are its MMA C fragments) */
    Q4K_DUMP_BULK        = 2,    /* gemv_q4k_tma_kernel, default float order */
    Q4K_DUMP_BULK_EXACT  = 3,    /* gemv_q4k_tma_kernel in the reference's AVX2 order */
    Q4K_DUMP_BIG         = 4,    /* gemv_q4k_tma_kernel in the tensor-pipe association: the large-matrix launch, N == 1 */
    Q4K_DUMP_WIDE        = 6,    /* gemv_q4k_wide_kernel: the batched decode kernel (8 < N <= 64 by default; any N here, 32 columns per launch) */
    Q4K_DUMP_GEMM        = 5,    /* gemm_q4k_tc_kernel: sumi = 8 hi + lo and summ as the epilogue reads them from TMEM; p_dev unused
                                  * (the tensor core never forms p[j]: the scales are folded into the A operand), N > 8 */
};
int q4k_mul_mat_dump(int kernel, const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                     const float *x_dev, int64_t N, int64_t x_row_stride_bytes, float *dst_dev,
                     int32_t *p_dev, int32_t *sumi_dev, int32_t *summ_dev,
                     void *workspace_dev, size_t workspace_bytes, void *stream);

/* Development / test knobs.  None of them is process-global mutable state: the defaults are read from the environment once
 * (Q4K_GEMV_ORDER, Q4K_DBG_*), and every setter below changes the CALLING THREAD's copy only, so backend instances driven by
 * different threads cannot change each other's numerics (SURVEY.md §8b "Threading"); the ggml plugin never calls a setter.
 * q4k_debug_set(name, value): "gemv_order", "fuse_limit", "gemm", "big_matrix_bytes", "pdl", "gemv_config", "mma_off",
 * "mma_kparts", "mma_stages", "mma_per_sm", "mma_contig", "tma_stages", "tma_per_sm", "gemm_ksplit", "midn_max",
 * "q6k_fuse_limit", "q6k_stages" (0 = the launch heuristic chooses), "q6k_gemm_min"; Q4K_ERR_UNSUPPORTED for an unknown name.
 * q4k_debug_reset(): this thread back to the process defaults. */
int  q4k_debug_set(const char *name, int64_t value);
void q4k_debug_reset(void);

/* Float order of the decode path (N <= 8): 0 / 2 = default — exact integer sums, fixed launch-independent float
 * association, results within 1e-5 of ggml_vec_dot_q4_K_q8_K; 1 = the reference's AVX2 accumulation order, fp32 results
 * bit-identical to the x86 build (slower kernel).  Overrides the Q4K_GEMV_ORDER environment variable ("exact" / "fast"). */
int q4k_debug_set_gemv_order(int mode);
/* development knob: 0 = launch heuristic, 1.. = force a compiled GEMV tile shape (gemv_q4k.cu) */
int q4k_debug_set_gemv_config(int cfg);

/* development knob: largest N*K (floats) for which q4k_mul_mat fuses quantize_row_q8_K into the GEMV launch
 * (default 16384); 0 = always run the quantizer as its own kernel. */
int q4k_debug_set_fuse_limit(int64_t n_floats);
/* development knob: 1 (default) = GEMV launches use programmatic dependent launch, 0 = plain stream order */
int q4k_debug_set_pdl(int on);

/* development knob: single-column (N == 1) launches whose weights are at least this many bytes (default 34 MiB) run
 * the lane-per-superblock bulk-copy kernel instead of the tensor-pipe kernel — same float association, identical bits,
 * faster where the matrix is large enough to be bound by the copy engine; <= 0 = never. */
int q4k_debug_set_big_matrix_bytes(int64_t bytes);

/* development knob: 1 (default) = N > 8 runs the tcgen05 GEMM, 0 = the decode kernel eight columns at a time */
int q4k_debug_set_gemm(int on);

/* host-buffer convenience used by the plugin's e2e path and bench.py's e2e leg: copies x (pinned or
 * pageable host) to the device, runs q4k_mul_mat, copies dst back.  Weights stay resident (w_dev).
 * Synchronous. */
int q4k_mul_mat_host(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                     const float *x_host, int64_t N, float *dst_host,
                     void *x_stage_dev, float *dst_stage_dev,
                     void *workspace_dev, size_t workspace_bytes, void *stream);

/* same without the final synchronize: everything (H2D of x, quantize, mul_mat, D2H of dst) is queued on
 * `stream`; the caller synchronizes the stream once per batch of calls.  This is what the backend plugin's
 * set_tensor_async / graph_compute / get_tensor_async + synchronize sequence amounts to
 * (L/ggml/src/ggml-backend-impl.h:95-106).  x_host / dst_host should be pinned. */
int q4k_mul_mat_host_async(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                           const float *x_host, int64_t N, float *dst_host,
                           void *x_stage_dev, float *dst_stage_dev,
                           void *workspace_dev, size_t workspace_bytes, void *stream);

/* ---- Q6_K weights ----------------------------------------------------------------------
 * TinyLlama Q4_K_M keeps attn_v / ffn_down of 10 layers and output.weight in Q6_K (L/src/llama-quant.cpp:228-230,
 * 302-303, 358-364); these entries replace ggml_compute_forward_mul_mat for Q6_K src0 / F32 src1
 * (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401 with type_traits_cpu[GGML_TYPE_Q6_K], :296-301: vec_dot =
 * ggml_vec_dot_q6_K_q8_K, L/ggml/src/ggml-cpu/arch/x86/quants.c:2129-2218; vec_dot_type = Q8_K).
 * Weights: M rows of K/256 block_q6_K (210 B each); w_dev and w_row_stride_bytes need only be even (210-byte blocks
 * are 2-byte aligned by nature).  N <= 8: the HBM-bound decode kernel; N > 8 ("q6k_gemm_min"): the
 * tcgen05 kind::f16 prefill GEMM (csrc/gemm_q6k_tc.cu; workspace 16-byte aligned).  Integer sums per superblock are exactly the
 * reference's on both; fp32 within 1e-5. */
size_t q6k_mul_mat_workspace_bytes(int64_t K, int64_t N);
int q6k_mul_mat(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                const float *x_dev, int64_t N, int64_t x_row_stride_bytes,
                float *dst_dev, int64_t dst_col_stride_bytes,
                void *workspace_dev, size_t workspace_bytes, void *stream);
/* src1 already Q8_K */
int q6k_mul_mat_q8(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                   const void *xq_dev, int64_t N, int64_t xq_row_stride_bytes,
                   float *dst_dev, int64_t dst_col_stride_bytes, void *stream);
/* parity: runs the PRODUCT kernel with its integer dump switched on: sumi_dev[(n*M + m)*(K/256) + b] = the int32
 * sum_j scales[j] * sum_l (q6 - 32) * q8 of superblock b (the reference's aux32 total, ggml-cpu/quants.c:740-752);
 * dst_dev = dense [N][M] fp32 results. */
int q6k_debug_partials(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes,
                       const void *xq_dev, int64_t N, int64_t xq_row_stride_bytes,
                       int32_t *sumi_dev, float *dst_dev, void *stream);

/* parity: the prefill GEMM's dump instantiation at any N >= 1 (F32 activations; workspace >= q6k_gemm_workspace_bytes, 16-byte
 * aligned): sumi_dev as above = the integers its epilogue read from tensor memory (16 * S_hi + S_lo, gemm_q6k_tc.cu). */
size_t q6k_gemm_workspace_bytes(int64_t K, int64_t N);
int q6k_mul_mat_dump(const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes, const float *x_dev, int64_t N,
                     int64_t x_row_stride_bytes, int32_t *sumi_dev, float *dst_dev, void *workspace_dev, size_t workspace_bytes,
                     void *stream);

/* ---- adjacent cheap ops of a decoder layer ----------------------------------------------
 * The F32 ops between the weight mul_mats of the reference's llama graph (build_norm / build_ffn / the residual adds,
 * L/src/llama-graph.cpp), so that the plugin keeps a norm -> mul_mat -> glu -> mul_mat -> add chain on the device and the
 * scheduler (L/ggml/src/ggml-backend.cpp:776-827) stops cutting the graph around every mul_mat.  Tensors are ggml's: dim 0 is
 * dense (4-byte elements), *_nb = byte strides of dims 1, 2, 3 (multiples of 4), ne = the four extents.  csrc/ops.cu. */
enum { Q4K_OP_ADD = 0, Q4K_OP_MUL = 1 };
/* GGML_OP_RMS_NORM, ggml_compute_forward_rms_norm_f32 (L/ggml/src/ggml-cpu/ops.cpp:3517-3570): y = x / sqrt(mean(x^2) + eps)
 * per row.  Optional fusions of the nodes around it in the llama graph (each NULL = off; results equal the separate ops bit for
 * bit, they are the same IEEE operations in the same order):
 *   res_dev: x := x + res first (the residual GGML_OP_ADD in front of the norm); sum_dev (optional) receives x + res;
 *   w_dev:   y := y * w, w = ne[0] floats broadcast over the rows (the GGML_OP_MUL by the norm weight behind the norm). */
int q4k_op_rms_norm(const float *x_dev, const int64_t *x_nb, const float *res_dev, const int64_t *res_nb, const float *w_dev,
                    float *y_dev, const int64_t *y_nb, float *sum_dev, const int64_t *sum_nb, const int64_t *ne, float eps, void *stream);
/* GGML_OP_ADD / GGML_OP_MUL, all F32: apply_binary_op (L/ggml/src/ggml-cpu/binary-ops.cpp:52-120); b is repeated over a
 * (ggml_can_repeat(b, a): every extent of a is a multiple of b's), dst has a's shape.  Bit-identical to the reference. */
int q4k_op_binary(int op, const float *a_dev, const int64_t *a_ne, const int64_t *a_nb, const float *b_dev, const int64_t *b_ne,
                  const int64_t *b_nb, float *dst_dev, const int64_t *dst_nb, void *stream);
/* GGML_OP_GLU / GGML_GLU_OP_SWIGLU, split form (ggml_swiglu_split: src0 = gate, src1 = up),
 * ggml_compute_forward_swiglu_f32 (L/ggml/src/ggml-cpu/ops.cpp:2934-2990): dst = silu(gate) * up, silu(x) = x / (1 + expf(-x))
 * (L/ggml/src/ggml-cpu/vec.h:957).  ~1e-6 relative against the reference's AVX2 polynomial expf. */
int q4k_op_swiglu(const float *gate_dev, int64_t gate_row_stride_bytes, const float *up_dev, int64_t up_row_stride_bytes,
                  float *dst_dev, int64_t dst_row_stride_bytes, int64_t nc, int64_t nrows, void *stream);
/* GGML_OP_GET_ROWS, F32 rows by I32 ids: ggml_compute_forward_get_rows_f32 (L/ggml/src/ggml-cpu/ops.cpp:4626-4660) */
int q4k_op_get_rows_f32(const float *src_dev, int64_t src_row_stride_bytes, int64_t nrows_src, const int32_t *ids_dev, int64_t n_ids,
                        float *dst_dev, int64_t dst_row_stride_bytes, int64_t nc, void *stream);
/* GGML_OP_ROPE, F32, modes NORMAL (neox = 0: pairs (2i, 2i+1)) and NEOX (pairs (i, i + n_dims/2)); no mrope / vision:
 * ggml_compute_forward_rope_f32 + ggml_rope_cache_init + rope_yarn (L/ggml/src/ggml-cpu/ops.cpp:5436-5475, 5534-5700).
 * x / dst: ne = [head_dim, heads, tokens, batch]; pos_dev: int32 per token; freq_factors_dev: optional n_dims/2 floats.
 * The caller hands over what the reference derives on the host: theta_scale = powf(freq_base, -2.0f / n_dims) and the YaRN
 * correction dims (ggml_rope_yarn_corr_dims, L/ggml/src/ggml.c:4122-4135).  The angle is built with the reference's own
 * sequence of float products; agreement with the reference is a few ulp of the rotated values (cosf / sinf, FMA contraction). */
int q4k_op_rope(const float *x_dev, const int64_t *x_nb, const int32_t *pos_dev, const float *freq_factors_dev, float *dst_dev,
                const int64_t *dst_nb, const int64_t *ne, int n_dims, int neox, float theta_scale, float freq_scale, float ext_factor,
                float attn_factor, float corr_lo, float corr_hi, void *stream);
/* ---- row-sharded weights over the GPUs of one process (SURVEY.md §8e inside the boundary) ---------------------------------
 * What ggml-cuda's split buffer does for --split-mode row (L/ggml/src/ggml-cuda/ggml-cuda.cu:774-787 get_row_split, :795-1000
 * the split buffer type / buffer, :1672 and :1720 the src1 / dst peer copies), for this path: a weight matrix is dealt to the
 * devices in contiguous row ranges (multiples of row_granule; fractions[i] = share of device i, NULL or all zero = equal) when it
 * is uploaded; q4k_split_mul_mat launches every device's shard on that device's own stream — the shard kernels read the
 * activations from the main device's memory and write their rows straight into the main device's dst over NVLink peer
 * mappings — and joins them into main_stream with events.  x_dev / dst_dev live on main_device; dst = N columns of M floats.
 * Results equal the single-GPU results bit for bit.  The plugin's split buffer type ("ggml_backend_split_buffer_type",
 * looked up by L/src/llama-model.cpp:382-394 for -sm row) sits on these.  Not thread-safe per device: shards of all split
 * matrices share one stream and one workspace per device. */
typedef struct q4k_split q4k_split;
q4k_split *q4k_split_create(const int *devices, int n_devices, const float *fractions, int64_t M, int64_t row_bytes, int row_granule);
void q4k_split_destroy(q4k_split *s);
int  q4k_split_parts(const q4k_split *s);
int  q4k_split_range(const q4k_split *s, int part, int *device, int64_t *row_low, int64_t *row_high);
/* buffer_i.set_tensor / get_tensor on a split tensor: bytes [offset, offset + size) of the full row-major matrix; synchronous */
int  q4k_split_set(q4k_split *s, const void *host, size_t offset, size_t size);
int  q4k_split_get(q4k_split *s, void *host, size_t offset, size_t size);
int  q4k_split_mul_mat(q4k_split *s, int q6k, int64_t K, const float *x_dev, int64_t N, int64_t x_row_stride_bytes, float *dst_dev,
                       int64_t dst_col_stride_bytes, int main_device, void *main_stream);

#ifdef __cplusplus
}
#endif
#endif /* Q4K_B200_H */
