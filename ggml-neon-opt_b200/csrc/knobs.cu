// knobs.cu — see knobs.h.  The q4k_debug_* entries of include/q4k_b200.h live here.
#include "knobs.h"

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <set>
#include <utility>

#include "../../include/q4k_b200.h"

namespace q4k {

namespace {

int env_int(const char *name, int dflt) {
    const char *e = getenv(name);
    return e ? atoi(e) : dflt;
}
int64_t env_i64(const char *name, int64_t dflt) {
    const char *e = getenv(name);
    return e ? atoll(e) : dflt;
}

Knobs read_defaults() {
    Knobs k;
    const char *ord = getenv("Q4K_GEMV_ORDER");
    k.order_mode = !ord ? 0 : (!strcmp(ord, "exact") ? 1 : (!strcmp(ord, "fast") ? 2 : 0));
    k.fuse_limit = env_i64("Q4K_DBG_FUSE_LIMIT", 16384);        // measured: K = 28672 is faster with the separate quantizer
    k.gemm_enabled = 1;
    k.big_matrix_bytes = env_i64("Q4K_DBG_BIG_BYTES", (int64_t)34 << 20);   // crossover measured: profiles/r1_big_sweep.txt
    k.pdl = env_int("Q4K_DBG_PDL", 1);
    k.gemv_config = 0;
    k.mma_off = env_int("Q4K_DBG_MMA", 0) < 0 ? 1 : 0;
    k.mma_kparts = env_int("Q4K_DBG_MMA_KPARTS", 0);
    k.mma_stages = env_int("Q4K_DBG_MMA_STAGES", 0);
    k.mma_per_sm = env_int("Q4K_DBG_MMA_PER_SM", 0);
    k.mma_contig = env_int("Q4K_DBG_MMA_CONTIG", 0);
    k.tma_stages = env_int("Q4K_DBG_STAGES", 0);
    k.tma_per_sm = env_int("Q4K_DBG_PER_SM", 0);
    k.gemm_ksplit = env_int("Q4K_DBG_GEMM_KSPLIT", 0);
    k.q6k_gemm_min = env_int("Q4K_DBG_Q6K_GEMM_MIN", 9);
    k.midn_max = env_int("Q4K_DBG_MIDN_MAX", 32);        // above 32 columns the tcgen05 GEMM wins on the large shapes (profiles/r2_wide_sweep.txt)
    k.wide_min = env_int("Q4K_DBG_WIDE_MIN", 2);       // measured: the batched kernel wins from N = 3 at every shape, at N = 2 from K = 4096 (profiles/r2_wide_sweep.txt)
    k.wide_dbg = env_int("Q4K_DBG_WIDE_DBG", 0);
    k.tma_boxes = env_int("Q4K_DBG_TMA_BOXES", 1);
    k.n1s_min_bytes = env_i64("Q4K_DBG_N1S_MIN_BYTES", 0);           // 0 = no size rule (development: 1 = always)
    k.n1s_min_k = env_i64("Q4K_DBG_N1S_MIN_K", 14336);
    k.n1s_grid = env_i64("Q4K_DBG_N1S_GRID", 0);
    k.grid_balance = env_int("Q4K_DBG_GRID_BALANCE", 3);
    k.n1s_ll_min_rows = env_i64("Q4K_DBG_N1S_LL_MIN_ROWS", 4096);
    k.q6k_fuse_limit = env_i64("Q6K_DBG_FUSE_LIMIT", 32768);
    k.q6k_stages = env_int("Q6K_DBG_STAGES", 0);
    return k;
}

const Knobs &process_defaults() {
    static const Knobs d = read_defaults();      // thread-safe one-time initialisation; never written again
    return d;
}

}  // namespace

Knobs &knobs() {
    thread_local Knobs t = process_defaults();
    return t;
}

void reset_knobs() { knobs() = process_defaults(); }

cudaError_t ensure_max_dynamic_smem(const void *kernel, int bytes) {
    static std::mutex mu;
    static std::set<std::pair<int, const void *>> done;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    const auto key = std::make_pair(dev, kernel);
    if (done.count(key)) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) done.insert(key);
    return e;
}

}  // namespace q4k

using q4k::knobs;

extern "C" {

int q4k_debug_set(const char *name, int64_t value) {
    if (!name) return Q4K_ERR_BAD_POINTER;
    q4k::Knobs &k = knobs();
    struct { const char *n; int *i; int64_t *l; } tab[] = {
        {"gemv_order", &k.order_mode, nullptr}, {"fuse_limit", nullptr, &k.fuse_limit}, {"gemm", &k.gemm_enabled, nullptr},
        {"big_matrix_bytes", nullptr, &k.big_matrix_bytes}, {"pdl", &k.pdl, nullptr}, {"gemv_config", &k.gemv_config, nullptr},
        {"mma_off", &k.mma_off, nullptr}, {"mma_kparts", &k.mma_kparts, nullptr}, {"mma_stages", &k.mma_stages, nullptr},
        {"mma_per_sm", &k.mma_per_sm, nullptr}, {"mma_contig", &k.mma_contig, nullptr}, {"tma_stages", &k.tma_stages, nullptr},
        {"tma_per_sm", &k.tma_per_sm, nullptr}, {"gemm_ksplit", &k.gemm_ksplit, nullptr}, {"midn_max", &k.midn_max, nullptr}, {"grid_balance", &k.grid_balance, nullptr}, {"q6k_gemm_min", &k.q6k_gemm_min, nullptr}, {"wide_min", &k.wide_min, nullptr}, {"wide_dbg", &k.wide_dbg, nullptr}, {"tma_boxes", &k.tma_boxes, nullptr}, {"n1s_min_bytes", nullptr, &k.n1s_min_bytes}, {"n1s_min_k", nullptr, &k.n1s_min_k}, {"n1s_grid", nullptr, &k.n1s_grid}, {"n1s_ll_min_rows", nullptr, &k.n1s_ll_min_rows},
        {"q6k_fuse_limit", nullptr, &k.q6k_fuse_limit}, {"q6k_stages", &k.q6k_stages, nullptr},
    };
    for (auto &t : tab) {
        if (strcmp(t.n, name)) continue;
        if (t.i) *t.i = (int)value; else *t.l = value;
        return Q4K_OK;
    }
    return Q4K_ERR_UNSUPPORTED;
}

void q4k_debug_reset(void) { q4k::reset_knobs(); }

int q4k_debug_set_gemv_order(int mode) {
    if (mode < 0 || mode > 2) return Q4K_ERR_UNSUPPORTED;
    knobs().order_mode = mode;
    return Q4K_OK;
}
int q4k_debug_set_gemv_config(int cfg) { knobs().gemv_config = cfg; return Q4K_OK; }
int q4k_debug_set_fuse_limit(int64_t n_floats) { knobs().fuse_limit = n_floats; return Q4K_OK; }
int q4k_debug_set_gemm(int on) { knobs().gemm_enabled = on ? 1 : 0; return Q4K_OK; }
int q4k_debug_set_big_matrix_bytes(int64_t bytes) { knobs().big_matrix_bytes = bytes; return Q4K_OK; }
int q4k_debug_set_pdl(int on) { knobs().pdl = on ? 1 : 0; return Q4K_OK; }

}  // extern "C"
