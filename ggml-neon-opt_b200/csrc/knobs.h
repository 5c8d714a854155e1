// knobs.h — development / test settings of the library, and per-device kernel attributes.
//
// Nothing here is process-global mutable state: the defaults are read from the environment ONCE (first use, then
// immutable); q4k_debug_set* change the CALLING THREAD's copy only.  A ggml backend instance is driven by one thread at a
// time (SURVEY.md §8b "Threading"), the plugin never calls a setter, so two backend instances / llama contexts cannot change
// each other's numerics, and no launch path calls getenv.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace q4k {

struct Knobs {
    int     order_mode;         // Q4K_GEMV_ORDER: 0 = default (tolerance order), 1 = the reference's AVX2 order bit for bit, 2 = never exact
    int64_t fuse_limit;         // N*K floats up to which quantize_row_q8_K runs inside the decode launch
    int     gemm_enabled;       // 0: N > 8 takes the decode kernel eight columns at a time
    int64_t big_matrix_bytes;   // N == 1 launches with at least this many weight bytes take the bulk-copy kernel (ORD 2)
    int     pdl;                // 1 = programmatic dependent launch
    int     gemv_config;        // gemv_q4k.cu tile shapes (development)
    int     mma_off;            // Q4K_DBG_MMA < 0: never the tensor-pipe decode kernel
    int     mma_kparts, mma_stages, mma_per_sm, mma_contig;   // gemv_q4k_mma.cu launch shape (0 = choose)
    int     tma_stages, tma_per_sm;                            // gemv_q4k_tma.cu
    int     gemm_ksplit;        // gemm_q4k_tc.cu: k > 0 forces a persistent grid of tiles x k CTAs; gemm_q6k_tc.cu: CTAs along K per tile (0 = choose)
    int     wide_min;           // smallest N q4k_mul_mat hands to the batched decode kernel (gemv_q4k_wide.cu); below it: gemv_q4k_mma.cu
    int64_t n1s_min_k;          // N == 1: matrices with at least this many columns (and >= 2048 rows) take the streamed kernel (gemv_q4k_n1s.cu)
    int64_t n1s_ll_min_rows;    // row-sharded LL form: shards with at least this many rows (and K >= n1s_min_k) take the streamed kernel
    int64_t n1s_grid;           // gemv_q4k_n1s.cu: CTAs of the launch (0 = choose)
    int     grid_balance;       // bit 0: n1s, bit 1: wide, bit 2: mma — persistent grids sized so that every CTA gets the same number of units
    int64_t n1s_min_bytes;      // ... or, if > 0, with at least this many weight bytes (development)
    int     tma_boxes;          // 1 = the decode kernels fetch weight tiles with TMA tensor copies (one per tile), 0 = one bulk copy per row
    int     wide_dbg;           // gemv_q4k_wide.cu timing experiments (results are wrong when set)
    int     midn_max;           // largest N served by the batched tensor-pipe decode kernel (gemv_q4k_wide.cu); 0 = off
    int64_t q6k_fuse_limit;
    int     q6k_stages;
    int     q6k_gemm_min;       // smallest N q6k_mul_mat hands to the tcgen05 GEMM (gemm_q6k_tc.cu); 0 = never
};

// A persistent grid of at most gmax CTAs over `units` equal work units: the smallest grid that needs no more rounds than gmax
// would — every CTA then works the same number of rounds, and no round runs with a fraction of the machine.
// Only for a few rounds: with many, the partial last round is a small share of the time and the CTAs given up for the balance
// cost more (measured, tools/knob_sweep.py: 8192 x 28672 N = 1, 256 passes: 148 CTAs 26.8 us, 128 CTAs 24.1 us; 28672 x 8192 N = 8,
// 448 passes: 28.7 -> 27.5 us; 28672 x 8192 N = 16, 896 passes: 43.6 -> 44.4 us, left alone).
inline int64_t balanced_grid(int64_t units, int64_t gmax) {
    if (units <= gmax || gmax < 1) return units;
    const int64_t rounds = (units + gmax - 1) / gmax;
    return rounds > 4 ? gmax : (units + rounds - 1) / rounds;
}

Knobs &knobs();                  // the calling thread's settings (initialised from the process defaults)
void reset_knobs();              // back to the process defaults (this thread)

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute of a kernel: remembered per (device, kernel), so a
// process that drives several GPUs (the plugin registers one device per sm_100 GPU) sets it on each of them.
cudaError_t ensure_max_dynamic_smem(const void *kernel, int bytes);

}  // namespace q4k
