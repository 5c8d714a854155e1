// ggml-q4k-b200.cpp — out-of-tree ggml backend plugin ("Q4K_B200") for NVIDIA B200.
//
// Claims exactly one operation: GGML_OP_MUL_MAT with GGML_TYPE_Q4_K src0, F32 src1, F32 dst — the path
// ggml_compute_forward_mul_mat -> quantize_row_q8_K -> ggml_vec_dot_q4_K_q8_K of the CPU backend
// (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401) — and, since Q4_K_M model files keep some matrices in it (SURVEY.md §8f
// rank 1), the same operation with GGML_TYPE_Q6_K src0 (ggml_vec_dot_q6_K_q8_K).  Everything else stays on whichever
// backend the scheduler picks.
// There is no fallback to ggml-cpu or ggml-cuda in here: a node this backend accepted either runs on the
// sm_100a kernels behind include/q4k_b200.h or graph_compute returns GGML_STATUS_FAILED.
//
// Loaded through the reference's own mechanism: ggml_backend_load(path) / env GGML_BACKEND_PATH
// (L/ggml/src/ggml-backend-reg.cpp:247-291, 603-607) -> ggml_backend_score() -> ggml_backend_init().
// Implements the vtables of L/ggml/src/ggml-backend-impl.h (API version 2): reg_i (:198-208), device_i
// (:140-184), buffer_type_i (:17-35), buffer_i (:41-58), backend_i (:90-126).  Shape of a one-op backend
// follows the role ggml-blas plays in the tree (L/ggml/src/ggml-blas/ggml-blas.cpp:227-518); none of its code
// is reused.
//
// This file is plain C++ (no CUDA headers): device memory, copies, streams and the kernels are reached
// through the C ABI of libq4k_b200.so only.
#include "ggml-backend-impl.h"
#include "ggml-impl.h"
#include "ggml.h"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/q4k_b200.h"

namespace {

constexpr size_t Q4K_TENSOR_ALIGN = 128;    // every tensor starts 128-byte aligned: 16 B loads / bulk copies are safe

struct reg_context;

struct device_context {
    int         cuda_device = 0;
    std::string name;           // "Q4K_B200_0"
    std::string description;    // "NVIDIA B200 (148 SMs)"
    ggml_backend_buffer_type buft{};
    ggml_backend_buffer_type host_buft{};   // page-locked host memory: what set_tensor_async / get_tensor_async copy from / to at full speed
};

struct buffer_context {
    int    cuda_device;
    void * base;
};

// A launch sequence captured once and replayed (CUDA graph with programmatic-dependent-launch edges).  `sig` identifies
// what was captured: every launching node's device pointers, shapes and strides.  ggml rebuilds the tensor structs of a graph
// for every token, but ggml-alloc hands the same addresses to the same topology, so a decode loop hits the cache.
struct graph_cache {
    std::vector<uint64_t> sig;
    q4k_graph * graph = nullptr;
    uint64_t ws_generation = 0;      // the captured launches hold the workspace pointer of this generation
    int seen = 0;                    // computes of `sig` so far (capture happens on the second: a graph seen once may never return)
    bool failed = false;             // capture of `sig` failed once: stay on plain launches
    ~graph_cache() { q4k_graph_destroy(graph); }
};

struct backend_context {
    device_context * dev;
    void *  stream    = nullptr;
    void *  workspace = nullptr;     // Q8_K copy of src1: the analogue of params->wdata (ggml-cpu.c:1267-1307)
    size_t  workspace_bytes = 0;
    uint64_t ws_generation = 1;      // bumped whenever the workspace is reallocated (invalidates captured launches)
    bool capturing = false;
    graph_cache cache;               // graph_compute's own cache; graph plans carry theirs
};

// backend_i.graph_plan_create's result (L/ggml/src/ggml-backend-impl.h:100-108): the node list + its captured launches
struct graph_plan {
    std::vector<ggml_tensor *> nodes;
    ggml_cgraph cgraph{};
    graph_cache cache;
};

std::vector<ggml_backend_device> & devices() {
    static std::vector<ggml_backend_device> d;
    return d;
}
std::vector<device_context> & device_contexts() {
    static std::vector<device_context> d;
    return d;
}

// ------------------------------------------------------------------------------------------ buffer
void buffer_free(ggml_backend_buffer_t buffer) {
    auto * ctx = (buffer_context *) buffer->context;
    q4k_mem_free(ctx->cuda_device, ctx->base);
    delete ctx;
}

void * buffer_get_base(ggml_backend_buffer_t buffer) {
    return ((buffer_context *) buffer->context)->base;
}

void buffer_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, uint8_t value, size_t offset, size_t size) {
    auto * ctx = (buffer_context *) buffer->context;
    if (q4k_mem_set(ctx->cuda_device, (char *) tensor->data + offset, value, size, nullptr) != Q4K_OK) {
        GGML_ABORT("Q4K_B200: memset_tensor failed (cudaError %d)", q4k_last_cuda_error());
    }
}

void buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    auto * ctx = (buffer_context *) buffer->context;
    if (q4k_mem_h2d(ctx->cuda_device, (char *) tensor->data + offset, data, size, nullptr) != Q4K_OK) {
        GGML_ABORT("Q4K_B200: set_tensor failed (cudaError %d)", q4k_last_cuda_error());
    }
}

void buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    auto * ctx = (buffer_context *) buffer->context;
    if (q4k_mem_d2h(ctx->cuda_device, data, (const char *) tensor->data + offset, size, nullptr) != Q4K_OK) {
        GGML_ABORT("Q4K_B200: get_tensor failed (cudaError %d)", q4k_last_cuda_error());
    }
}

bool buffer_cpy_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * src, ggml_tensor * dst) {
    auto * ctx = (buffer_context *) buffer->context;
    if (!src->buffer) return false;
    if (ggml_backend_buffer_is_host(src->buffer)) {
        return q4k_mem_h2d(ctx->cuda_device, dst->data, src->data, ggml_nbytes(src), nullptr) == Q4K_OK;
    }
    if (src->buffer->iface.get_base == buffer_get_base &&
        ((buffer_context *) src->buffer->context)->cuda_device == ctx->cuda_device) {
        return q4k_mem_d2d(ctx->cuda_device, dst->data, src->data, ggml_nbytes(src), nullptr) == Q4K_OK;
    }
    return false;       // the caller falls back to get_tensor + set_tensor through host memory
}

void buffer_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    auto * ctx = (buffer_context *) buffer->context;
    if (q4k_mem_set(ctx->cuda_device, ctx->base, value, buffer->size, nullptr) != Q4K_OK) {
        GGML_ABORT("Q4K_B200: clear failed (cudaError %d)", q4k_last_cuda_error());
    }
}

const ggml_backend_buffer_i buffer_iface = {
    /* .free_buffer   = */ buffer_free,
    /* .get_base      = */ buffer_get_base,
    /* .init_tensor   = */ nullptr,
    /* .memset_tensor = */ buffer_memset_tensor,
    /* .set_tensor    = */ buffer_set_tensor,
    /* .get_tensor    = */ buffer_get_tensor,
    /* .cpy_tensor    = */ buffer_cpy_tensor,
    /* .clear         = */ buffer_clear,
    /* .reset         = */ nullptr,
};

// ------------------------------------------------------------------------------------------ buffer type
const char * buft_get_name(ggml_backend_buffer_type_t buft) {
    return ((device_context *) buft->context)->name.c_str();
}

ggml_backend_buffer_t buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    auto * dev = (device_context *) buft->context;
    void * base = q4k_mem_alloc(dev->cuda_device, size + Q4K_TENSOR_ALIGN);
    if (!base) {
        GGML_LOG_ERROR("Q4K_B200: failed to allocate %zu bytes on device %d (cudaError %d)\n", size, dev->cuda_device, q4k_last_cuda_error());
        return nullptr;                                   // the caller treats NULL as out-of-memory
    }
    auto * ctx = new buffer_context{dev->cuda_device, base};
    return ggml_backend_buffer_init(buft, buffer_iface, ctx, size);
}

size_t buft_get_alignment(ggml_backend_buffer_type_t) { return Q4K_TENSOR_ALIGN; }

const ggml_backend_buffer_type_i buft_iface = {
    /* .get_name       = */ buft_get_name,
    /* .alloc_buffer   = */ buft_alloc_buffer,
    /* .get_alignment  = */ buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ nullptr,
    /* .is_host        = */ nullptr,
};

bool buft_is_ours(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == buft_get_name; }

// ------------------------------------------------------------------------------------------ host (pinned) buffer type
// device_i.get_host_buffer_type (L/ggml/src/ggml-backend-impl.h:158): an ordinary CPU buffer whose pages are locked, so the
// asynchronous copies of the backend run at PCIe speed and really are asynchronous.  The buffer itself is the reference's own CPU
// buffer (ggml_backend_cpu_buffer_from_ptr); only the allocation and the free are ours.
const char * host_buft_get_name(ggml_backend_buffer_type_t) { return "Q4K_B200_Host"; }

void host_buffer_free(ggml_backend_buffer_t buffer) { q4k_host_free_pinned(buffer->context); }

ggml_backend_buffer_t host_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    void * ptr = q4k_host_alloc_pinned(size);
    if (!ptr) return nullptr;                                  // the caller falls back to a pageable CPU buffer
    ggml_backend_buffer_t buffer = ggml_backend_cpu_buffer_from_ptr(ptr, size);
    buffer->buft = buft;
    buffer->iface.free_buffer = host_buffer_free;
    return buffer;
}

size_t host_buft_get_alignment(ggml_backend_buffer_type_t) { return Q4K_TENSOR_ALIGN; }
bool host_buft_is_host(ggml_backend_buffer_type_t) { return true; }

const ggml_backend_buffer_type_i host_buft_iface = {
    /* .get_name       = */ host_buft_get_name,
    /* .alloc_buffer   = */ host_buft_alloc_buffer,
    /* .get_alignment  = */ host_buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ nullptr,
    /* .is_host        = */ host_buft_is_host,
};

// ------------------------------------------------------------------------------------------ split (row-sharded) buffer type
// --split-mode row: L/src/llama-model.cpp:375-401 asks the registry for "ggml_backend_split_buffer_type" and puts the matmul
// weights of a layer into the buffer type it returns.  The shape follows ggml-cuda's (L/ggml/src/ggml-cuda/ggml-cuda.cu:795-1050):
// the buffer has no memory of its own (fake base address), init_tensor deals the tensor's rows to the devices
// (q4k_split_create), set_tensor / get_tensor scatter / gather, and MUL_MAT on such a tensor runs every device's shard with the
// result rows stored straight into the main device's dst (q4k_split_mul_mat).  Everything below the vtable is the C ABI.
struct split_buft_context {
    int main_device = 0;                 // index into devices()
    std::vector<float> fractions;        // one share per registered device (all zero = equal)
    std::string name;
};

struct split_buffer_context {
    std::vector<q4k_split *> splits;     // one per tensor, owned
    ~split_buffer_context() { for (auto * s : splits) q4k_split_destroy(s); }
};

const char * split_buft_get_name(ggml_backend_buffer_type_t buft) { return ((split_buft_context *) buft->context)->name.c_str(); }
bool buft_is_split(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == split_buft_get_name; }
bool tensor_is_split(const ggml_tensor * t) { return t && t->buffer && buft_is_split(t->buffer->buft); }

void split_buffer_free(ggml_backend_buffer_t buffer) { delete (split_buffer_context *) buffer->context; }
void * split_buffer_get_base(ggml_backend_buffer_t) { return (void *) 0x1000; }      // never dereferenced: tensors carry their shards in ->extra

ggml_status split_buffer_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor) {
    GGML_ASSERT(tensor->view_src == nullptr && "views of split tensors are not supported");
    GGML_ASSERT(ggml_is_contiguous(tensor) && "split buffers hold contiguous tensors only");
    auto * ctx = (split_buffer_context *) buffer->context;
    auto * bctx = (split_buft_context *) buffer->buft->context;
    std::vector<int> devs;
    for (auto & d : device_contexts()) devs.push_back(d.cuda_device);
    // 16-row granule: the decode kernels' tile, so every shard starts on a tile boundary
    q4k_split * sp = q4k_split_create(devs.data(), (int) devs.size(), bctx->fractions.data(), ggml_nrows(tensor),
                                      (int64_t) ggml_row_size(tensor->type, tensor->ne[0]), 16);
    if (!sp) {
        GGML_LOG_ERROR("Q4K_B200: cannot allocate the shards of %s (cudaError %d)\n", tensor->name, q4k_last_cuda_error());
        return GGML_STATUS_ALLOC_FAILED;
    }
    ctx->splits.push_back(sp);
    tensor->extra = sp;
    return GGML_STATUS_SUCCESS;
}

void split_buffer_set_tensor(ggml_backend_buffer_t, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    if (q4k_split_set((q4k_split *) tensor->extra, data, offset, size) != Q4K_OK)
        GGML_ABORT("Q4K_B200: set_tensor on split tensor %s failed (cudaError %d)", tensor->name, q4k_last_cuda_error());
}

void split_buffer_get_tensor(ggml_backend_buffer_t, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    if (q4k_split_get((q4k_split *) tensor->extra, data, offset, size) != Q4K_OK)
        GGML_ABORT("Q4K_B200: get_tensor on split tensor %s failed (cudaError %d)", tensor->name, q4k_last_cuda_error());
}

void split_buffer_clear(ggml_backend_buffer_t, uint8_t) {}

const ggml_backend_buffer_i split_buffer_iface = {
    /* .free_buffer   = */ split_buffer_free,
    /* .get_base      = */ split_buffer_get_base,
    /* .init_tensor   = */ split_buffer_init_tensor,
    /* .memset_tensor = */ nullptr,
    /* .set_tensor    = */ split_buffer_set_tensor,
    /* .get_tensor    = */ split_buffer_get_tensor,
    /* .cpy_tensor    = */ nullptr,
    /* .clear         = */ split_buffer_clear,
    /* .reset         = */ nullptr,
};

ggml_backend_buffer_t split_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    return ggml_backend_buffer_init(buft, split_buffer_iface, new split_buffer_context, size);
}
size_t split_buft_get_alignment(ggml_backend_buffer_type_t) { return Q4K_TENSOR_ALIGN; }
size_t split_buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * tensor) { return ggml_nbytes(tensor); }
bool split_buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i split_buft_iface = {
    /* .get_name       = */ split_buft_get_name,
    /* .alloc_buffer   = */ split_buft_alloc_buffer,
    /* .get_alignment  = */ split_buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ split_buft_get_alloc_size,
    /* .is_host        = */ split_buft_is_host,
};

// the proc L/src/llama-model.cpp looks up (ggml_backend_split_buffer_type_t: (int main_device, const float * tensor_split));
// one buffer type object per distinct (main device, shares), kept for the life of the process like the registry itself
ggml_backend_buffer_type_t split_buffer_type(int main_device, const float * tensor_split) {
    static std::mutex mutex;
    static std::vector<ggml_backend_buffer_type *> made;
    std::lock_guard<std::mutex> lock(mutex);
    const size_t n = devices().size();
    if (n == 0 || main_device < 0 || (size_t) main_device >= n) return nullptr;
    std::vector<float> fr(n, 0.0f);
    if (tensor_split) for (size_t i = 0; i < n; ++i) fr[i] = tensor_split[i];
    for (auto * b : made) {
        auto * c = (split_buft_context *) b->context;
        if (c->main_device == main_device && c->fractions == fr) return b;
    }
    auto * c = new split_buft_context;
    c->main_device = main_device;
    c->fractions = fr;
    c->name = "Q4K_B200_Split";
    made.push_back(new ggml_backend_buffer_type{split_buft_iface, &devices()[main_device], c});
    return made.back();
}

}  // namespace

namespace {

// ------------------------------------------------------------------------------------------ backend (stream)
const char * backend_get_name(ggml_backend_t) { return "Q4K_B200"; }

void backend_free(ggml_backend_t backend) {
    auto * ctx = (backend_context *) backend->context;
    q4k_stream_synchronize(ctx->dev->cuda_device, ctx->stream);
    q4k_mem_free(ctx->dev->cuda_device, ctx->workspace);
    q4k_stream_destroy(ctx->dev->cuda_device, ctx->stream);
    delete ctx;
    delete backend;
}

void backend_set_tensor_async(ggml_backend_t backend, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    auto * ctx = (backend_context *) backend->context;
    if (q4k_mem_h2d(ctx->dev->cuda_device, (char *) tensor->data + offset, data, size, ctx->stream) != Q4K_OK) {
        GGML_ABORT("Q4K_B200: set_tensor_async failed (cudaError %d)", q4k_last_cuda_error());
    }
}

void backend_get_tensor_async(ggml_backend_t backend, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    auto * ctx = (backend_context *) backend->context;
    if (q4k_mem_d2h(ctx->dev->cuda_device, data, (const char *) tensor->data + offset, size, ctx->stream) != Q4K_OK) {
        GGML_ABORT("Q4K_B200: get_tensor_async failed (cudaError %d)", q4k_last_cuda_error());
    }
}

void backend_synchronize(ggml_backend_t backend) {
    auto * ctx = (backend_context *) backend->context;
    if (q4k_stream_synchronize(ctx->dev->cuda_device, ctx->stream) != Q4K_OK) {
        GGML_ABORT("Q4K_B200: synchronize failed (cudaError %d)", q4k_last_cuda_error());
    }
}

// the conditions under which this backend takes a MUL_MAT node (also the contract of supports_op)
bool mul_mat_supported(const ggml_tensor * op) {
    const ggml_tensor * src0 = op->src[0];
    const ggml_tensor * src1 = op->src[1];
    if (!src0 || !src1) return false;
    const bool q6 = src0->type == GGML_TYPE_Q6_K;
    if ((src0->type != GGML_TYPE_Q4_K && !q6) || src1->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32) return false;
    if (src0->ne[0] % Q4K_QK_K != 0 || src0->ne[0] != src1->ne[0]) return false;
    if (q6) {
        // weights: whole block_q6_K rows; 210-byte blocks are 2-byte aligned by nature, the kernel takes any even stride
        if (src0->nb[0] != Q4K_BLOCK_Q6K || src0->nb[1] < (size_t) (src0->ne[0] / Q4K_QK_K) * Q4K_BLOCK_Q6K) return false;
        if ((src0->nb[1] | src0->nb[2] | src0->nb[3]) & 1) return false;
    } else {
        // weights: whole block_q4_K rows, 16-byte aligned strides (144 = 9 x 16, so any stride made of rows is)
        if (src0->nb[0] != Q4K_BLOCK_Q4K || src0->nb[1] < (size_t) (src0->ne[0] / Q4K_QK_K) * Q4K_BLOCK_Q4K) return false;
        if ((src0->nb[1] | src0->nb[2] | src0->nb[3]) & 15) return false;
    }
    // activations: unit inner stride, 16-byte aligned row / batch strides (permuted src1 is declined)
    if (src1->nb[0] != sizeof(float) || src1->nb[1] < (size_t) src1->ne[0] * sizeof(float)) return false;
    if ((src1->nb[1] | src1->nb[2] | src1->nb[3]) & 15) return false;
    // result: dense columns
    if (op->nb[0] != sizeof(float) || op->nb[1] < (size_t) op->ne[0] * sizeof(float)) return false;
    if ((op->nb[1] | op->nb[2] | op->nb[3]) & 3) return false;
    // broadcast of src0 over dims 2 / 3 as in ggml_compute_forward_mul_mat (ggml-cpu.c:1142-1143)
    if (src0->ne[2] == 0 || src0->ne[3] == 0) return false;
    if (src1->ne[2] % src0->ne[2] != 0 || src1->ne[3] % src0->ne[3] != 0) return false;
    // the weights must live in this backend's memory: tensor->data is a device pointer here — or in its split buffer type
    // (rows dealt to the devices, shards in ->extra): whole 2-D matrices only, activations and result dense
    if (tensor_is_split(src0)) {
        return src0->ne[2] == 1 && src0->ne[3] == 1 && src1->ne[2] == 1 && src1->ne[3] == 1 && ggml_is_contiguous(src0) && !src0->view_src &&
               op->nb[1] >= (size_t) src0->ne[1] * sizeof(float);
    }
    if (src0->buffer && !buft_is_ours(src0->buffer->buft)) return false;
    return true;
}

ggml_status compute_mul_mat(backend_context * ctx, ggml_tensor * node, bool reuse_quantized) {
    const ggml_tensor * src0 = node->src[0];
    const ggml_tensor * src1 = node->src[1];
    const int64_t K = src0->ne[0], M = src0->ne[1], N = src1->ne[1];
    const int64_t ne02 = src0->ne[2], ne03 = src0->ne[3], ne12 = src1->ne[2], ne13 = src1->ne[3];
    const int64_t r2 = ne12 / ne02, r3 = ne13 / ne03;
    if (M == 0 || N == 0 || ne12 == 0 || ne13 == 0) return GGML_STATUS_SUCCESS;

    const bool q6 = src0->type == GGML_TYPE_Q6_K;
    if (tensor_is_split(src0)) {           // row-sharded weights: every device runs its rows, results land in this device's dst
        const int rc = q4k_split_mul_mat((q4k_split *) src0->extra, q6 ? 1 : 0, K, (const float *) src1->data, N, (int64_t) src1->nb[1],
                                         (float *) node->data, (int64_t) node->nb[1], ctx->dev->cuda_device, ctx->stream);
        if (rc != Q4K_OK) {
            GGML_LOG_ERROR("Q4K_B200: q4k_split_mul_mat(%s) failed: rc=%d cudaError=%d\n", node->name, rc, q4k_last_cuda_error());
            return GGML_STATUS_FAILED;
        }
        return GGML_STATUS_SUCCESS;
    }
    const size_t need = q6 ? q6k_mul_mat_workspace_bytes(K, N) : q4k_mul_mat_workspace_bytes(K, N);
    if (need > ctx->workspace_bytes) {
        GGML_LOG_ERROR("Q4K_B200: %s: workspace was not sized before the launches (need %zu, have %zu)\n", node->name, need, ctx->workspace_bytes);
        return GGML_STATUS_FAILED;
    }
    for (int64_t i13 = 0; i13 < ne13; ++i13) {
        for (int64_t i12 = 0; i12 < ne12; ++i12) {
            const int64_t i03 = i13 / r3, i02 = i12 / r2;
            const char * w = (const char *) src0->data + i02 * src0->nb[2] + i03 * src0->nb[3];
            const char * x = (const char *) src1->data + i12 * src1->nb[2] + i13 * src1->nb[3];
            char * d = (char *) node->data + i12 * node->nb[2] + i13 * node->nb[3];
            int rc = Q4K_ERR_UNSUPPORTED;
            if (reuse_quantized && !q6) {      // the node in front quantized this src1 into the workspace: GEMM only
                rc = q4k_mul_mat_reuse(w, M, K, (int64_t) src0->nb[1], N, (float *) d, (int64_t) node->nb[1], ctx->workspace,
                                       ctx->workspace_bytes, ctx->stream);
            }
            if (rc == Q4K_ERR_UNSUPPORTED)
            rc = q6 ? q6k_mul_mat(w, M, K, (int64_t) src0->nb[1], (const float *) x, N, (int64_t) src1->nb[1],
                                            (float *) d, (int64_t) node->nb[1], ctx->workspace, ctx->workspace_bytes, ctx->stream)
                              : q4k_mul_mat(w, M, K, (int64_t) src0->nb[1], (const float *) x, N, (int64_t) src1->nb[1],
                                            (float *) d, (int64_t) node->nb[1], ctx->workspace, ctx->workspace_bytes, ctx->stream);
            if (rc != Q4K_OK) {
                GGML_LOG_ERROR("Q4K_B200: %s(%s) failed: rc=%d cudaError=%d\n", q6 ? "q6k_mul_mat" : "q4k_mul_mat", node->name, rc, q4k_last_cuda_error());
                return GGML_STATUS_FAILED;
            }
        }
    }
    return GGML_STATUS_SUCCESS;
}

// A mul_mat that can share a launch with its neighbours: plain 2-D, dense rows on both sides (decode batches: one launch of the
// decode kernel; prefill batches: one GEMM launch over all row tiles; q4k_mul_mat_group_ws declines the sizes in between).
bool groupable(const ggml_tensor * node) {
    const ggml_tensor * src0 = node->src[0];
    const ggml_tensor * src1 = node->src[1];
    if (tensor_is_split(src0)) return false;
    return src0->type == GGML_TYPE_Q4_K && src0->ne[2] == 1 && src0->ne[3] == 1 && src1->ne[2] == 1 && src1->ne[3] == 1 && src1->ne[1] >= 1 &&
           src0->ne[1] > 0 && node->nb[1] == (size_t) src0->ne[1] * sizeof(float);
}

bool is_noop(const ggml_tensor * node);

// byte ranges of two tensors overlap (views carry their own data pointer; ggml_nbytes covers strided extents)
bool overlaps(const ggml_tensor * a, const ggml_tensor * b) {
    if (!a || !b || !a->data || !b->data) return false;
    const char * a0 = (const char *) a->data, * b0 = (const char *) b->data;
    return a0 < b0 + ggml_nbytes(b) && b0 < a0 + ggml_nbytes(a);
}

// May mul_mat node `h` run BEFORE the launching node `b` that precedes it in graph order?  (ggml-alloc reuses the memory of dead
// tensors for later nodes, so "no data dependency" has to be checked on addresses, not on the graph.)
bool can_hoist_over(const ggml_tensor * h, const ggml_tensor * b) {
    if (overlaps(h, b)) return false;                                        // b's result would land on h's
    for (int k = 0; k < GGML_MAX_SRC; ++k) {
        if (overlaps(h, b->src[k])) return false;                            // h's result would land on something b still reads
        if (overlaps(h->src[k], b)) return false;                            // h reads what b writes: a true dependency
    }
    return true;
}

// MUL_MAT nodes that read the same activations (attn_q / attn_k / attn_v, ffn_gate / ffn_up: separate ggml_mul_mat nodes on
// one `cur` in the reference's llama graph) go out as ONE launch (q4k_mul_mat_group).  The members need not be adjacent: the
// llama graph puts rope(q) between attn_q and attn_k (L/src/llama-graph.cpp build_attn: q_cur, k_cur, v_cur are expanded one
// after the other); a later member is pulled in front of such nodes when no byte range says otherwise (can_hoist_over).
// Marks every consumed node in `done`; returns the number of mul_mats fused (0 = not fused, -1 = launch failed).
int try_fused_group(backend_context * ctx, ggml_cgraph * cgraph, int i, std::vector<char> & done) {
    ggml_tensor * first = cgraph->nodes[i];
    if (!groupable(first)) return 0;
    const void * w[4];
    int64_t M[4];
    const float * x[4];
    float * dst[4];
    int idx[4];
    int n = 0;
    std::vector<const ggml_tensor *> skipped;                                 // launching nodes between the members
    for (int j = i; j < cgraph->n_nodes && n < 4 && j < i + 12; ++j) {
        ggml_tensor * node = cgraph->nodes[j];
        if (is_noop(node) || done[j]) continue;
        const bool member = node->op == GGML_OP_MUL_MAT && mul_mat_supported(node) && groupable(node) && node->src[1] == first->src[1] &&
                            node->src[0]->ne[0] == first->src[0]->ne[0] && node->src[0]->nb[1] == first->src[0]->nb[1];
        bool ok = member;
        for (size_t k = 0; ok && k < skipped.size(); ++k) ok = can_hoist_over(node, skipped[k]);
        for (int k = 0; ok && k < n; ++k) ok = !overlaps(node, cgraph->nodes[idx[k]]);     // two members never share a result range
        if (!ok) {
            if (node->op == GGML_OP_MUL_MAT) break;                           // a mul_mat that cannot join ends the search
            skipped.push_back(node);
            continue;
        }
        w[n] = node->src[0]->data; M[n] = node->src[0]->ne[1]; x[n] = (const float *) node->src[1]->data; dst[n] = (float *) node->data;
        idx[n] = j;
        ++n;
    }
    if (n < 2) return 0;
    const ggml_tensor * src1 = first->src[1];
    const int rc = q4k_mul_mat_group_ws(n, w, M, first->src[0]->ne[0], (int64_t) first->src[0]->nb[1], x[0], src1->ne[1], (int64_t) src1->nb[1], dst,
                                        ctx->workspace, ctx->workspace_bytes, ctx->stream);
    if (rc == Q4K_ERR_UNSUPPORTED) return 0;
    if (rc != Q4K_OK) {
        GGML_LOG_ERROR("Q4K_B200: q4k_mul_mat_group_ws(%s ...) failed: rc=%d cudaError=%d\n", first->name, rc, q4k_last_cuda_error());
        return -1;
    }
    for (int k = 0; k < n; ++k) done[idx[k]] = 1;
    return n;
}

// ---- the cheap F32 ops around the mul_mats (kernels: csrc/ops.cu) -------------------------------
bool f32_rows(const ggml_tensor * t) {       // F32, dense dim 0, 4-byte-multiple strides
    return t && t->type == GGML_TYPE_F32 && t->nb[0] == sizeof(float) && !((t->nb[1] | t->nb[2] | t->nb[3]) & 3);
}

bool rms_norm_supported(const ggml_tensor * op) {
    return f32_rows(op->src[0]) && f32_rows(op) && ggml_are_same_shape(op->src[0], op) && ggml_nrows(op) <= 0x7fffffff;
}

bool binary_supported(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0], * b = op->src[1];
    return f32_rows(a) && f32_rows(b) && f32_rows(op) && ggml_are_same_shape(a, op) && ggml_can_repeat(b, a) && ggml_nrows(op) <= 65535;
}

bool glu_supported(const ggml_tensor * op) {
    const ggml_tensor * g = op->src[0], * u = op->src[1];
    if (ggml_get_glu_op(op) != GGML_GLU_OP_SWIGLU || !u) return false;                 // split form only (what build_ffn emits)
    return f32_rows(g) && f32_rows(u) && f32_rows(op) && ggml_is_contiguous_1(g) && ggml_is_contiguous_1(u) && ggml_is_contiguous_1(op) &&
           ggml_are_same_shape(g, u) && op->ne[0] == g->ne[0] && ggml_nrows(op) == ggml_nrows(g) && ggml_nrows(op) <= 65535;
}

bool get_rows_supported(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0], * ids = op->src[1];
    return f32_rows(a) && f32_rows(op) && ids && ids->type == GGML_TYPE_I32 && ggml_is_contiguous(ids) && a->ne[2] == 1 && a->ne[3] == 1 &&
           ids->ne[1] == 1 && ids->ne[2] == 1 && ids->ne[3] == 1 && ids->ne[0] <= 65535;
}

bool rope_supported(const ggml_tensor * op) {
    const ggml_tensor * x = op->src[0], * pos = op->src[1], * ff = op->src[2];
    const int32_t * prm = (const int32_t *) op->op_params;
    const int n_dims = prm[1], mode = prm[2];
    if (mode != GGML_ROPE_TYPE_NORMAL && mode != GGML_ROPE_TYPE_NEOX) return false;            // no mrope / vision
    if (!f32_rows(x) || !f32_rows(op) || !ggml_are_same_shape(x, op) || !pos || pos->type != GGML_TYPE_I32 || !ggml_is_contiguous(pos)) return false;
    if (pos->ne[0] != x->ne[2] || (x->ne[0] & 1) || n_dims <= 0 || (n_dims & 1) || n_dims > x->ne[0]) return false;
    if (ff && (ff->type != GGML_TYPE_F32 || !ggml_is_contiguous(ff) || ff->ne[0] < n_dims / 2)) return false;
    return x->ne[1] <= 65535 && x->ne[2] * x->ne[3] <= 65535;
}

ggml_status op_failed(const char * what, const ggml_tensor * node, int rc) {
    GGML_LOG_ERROR("Q4K_B200: %s(%s) failed: rc=%d cudaError=%d\n", what, node->name, rc, q4k_last_cuda_error());
    return GGML_STATUS_FAILED;
}

// RMS_NORM, optionally with the residual ADD in front of it (`add`: its result is still written — the residual stream needs it)
// and the MUL by the norm weight behind it (`mul`) in the same launch
ggml_status compute_rms_norm(backend_context * ctx, const ggml_tensor * add, const ggml_tensor * norm, const ggml_tensor * mul) {
    const ggml_tensor * x = add ? add->src[0] : norm->src[0];
    const ggml_tensor * res = add ? add->src[1] : nullptr;
    const ggml_tensor * out = mul ? mul : norm;
    const ggml_tensor * w = mul ? (mul->src[0] == norm ? mul->src[1] : mul->src[0]) : nullptr;
    float eps;
    memcpy(&eps, norm->op_params, sizeof(float));
    const int64_t x_nb[3] = {(int64_t) x->nb[1], (int64_t) x->nb[2], (int64_t) x->nb[3]};
    const int64_t y_nb[3] = {(int64_t) out->nb[1], (int64_t) out->nb[2], (int64_t) out->nb[3]};
    int64_t r_nb[3] = {0, 0, 0}, s_nb[3] = {0, 0, 0};
    if (add) {
        for (int d = 0; d < 3; ++d) { r_nb[d] = (int64_t) res->nb[d + 1]; s_nb[d] = (int64_t) add->nb[d + 1]; }
    }
    const int rc = q4k_op_rms_norm((const float *) x->data, x_nb, res ? (const float *) res->data : nullptr, r_nb, w ? (const float *) w->data : nullptr,
                                   (float *) out->data, y_nb, add ? (float *) add->data : nullptr, s_nb, norm->ne, eps, ctx->stream);
    return rc == Q4K_OK ? GGML_STATUS_SUCCESS : op_failed("q4k_op_rms_norm", norm, rc);
}

// how many nodes from cgraph->nodes[i] go into one q4k_op_rms_norm launch: [ADD] RMS_NORM [MUL by a broadcast weight row]
int norm_chain(const ggml_cgraph * cgraph, int i, const ggml_tensor *& add, const ggml_tensor *& norm, const ggml_tensor *& mul) {
    add = norm = mul = nullptr;
    int j = i;
    const ggml_tensor * n0 = cgraph->nodes[j];
    if (n0->op == GGML_OP_ADD) {
        if (j + 1 >= cgraph->n_nodes || !binary_supported(n0) || !ggml_are_same_shape(n0->src[0], n0->src[1])) return 0;
        const ggml_tensor * n1 = cgraph->nodes[j + 1];
        if (n1->op != GGML_OP_RMS_NORM || n1->src[0] != n0 || !rms_norm_supported(n1)) return 0;
        add = n0;
        ++j;
    }
    if (cgraph->nodes[j]->op != GGML_OP_RMS_NORM || !rms_norm_supported(cgraph->nodes[j])) return 0;
    norm = cgraph->nodes[j];
    if (j + 1 < cgraph->n_nodes && ggml_can_fuse(cgraph, j, {GGML_OP_RMS_NORM, GGML_OP_MUL})) {
        const ggml_tensor * m = cgraph->nodes[j + 1];
        const ggml_tensor * w = m->src[0] == norm ? m->src[1] : m->src[0];
        if (binary_supported(m) && w != norm && f32_rows(w) && w->ne[0] == norm->ne[0] && ggml_nrows(w) == 1) { mul = m; ++j; }
    }
    return j - i + 1;
}

ggml_status compute_binary(backend_context * ctx, const ggml_tensor * node) {
    const ggml_tensor * a = node->src[0], * b = node->src[1];
    const int64_t a_nb[3] = {(int64_t) a->nb[1], (int64_t) a->nb[2], (int64_t) a->nb[3]};
    const int64_t b_nb[3] = {(int64_t) b->nb[1], (int64_t) b->nb[2], (int64_t) b->nb[3]};
    const int64_t d_nb[3] = {(int64_t) node->nb[1], (int64_t) node->nb[2], (int64_t) node->nb[3]};
    const int rc = q4k_op_binary(node->op == GGML_OP_ADD ? Q4K_OP_ADD : Q4K_OP_MUL, (const float *) a->data, a->ne, a_nb, (const float *) b->data, b->ne,
                                 b_nb, (float *) node->data, d_nb, ctx->stream);
    return rc == Q4K_OK ? GGML_STATUS_SUCCESS : op_failed("q4k_op_binary", node, rc);
}

ggml_status compute_glu(backend_context * ctx, const ggml_tensor * node) {
    const ggml_tensor * g = node->src[0], * u = node->src[1];
    const int rc = q4k_op_swiglu((const float *) g->data, (int64_t) g->nb[1], (const float *) u->data, (int64_t) u->nb[1], (float *) node->data,
                                 (int64_t) node->nb[1], node->ne[0], ggml_nrows(node), ctx->stream);
    return rc == Q4K_OK ? GGML_STATUS_SUCCESS : op_failed("q4k_op_swiglu", node, rc);
}

ggml_status compute_get_rows(backend_context * ctx, const ggml_tensor * node) {
    const ggml_tensor * a = node->src[0], * ids = node->src[1];
    const int rc = q4k_op_get_rows_f32((const float *) a->data, (int64_t) a->nb[1], a->ne[1], (const int32_t *) ids->data, ids->ne[0],
                                       (float *) node->data, (int64_t) node->nb[1], node->ne[0], ctx->stream);
    return rc == Q4K_OK ? GGML_STATUS_SUCCESS : op_failed("q4k_op_get_rows_f32", node, rc);
}

ggml_status compute_rope(backend_context * ctx, const ggml_tensor * node) {
    const ggml_tensor * x = node->src[0], * pos = node->src[1], * ff = node->src[2];
    const int32_t * prm = (const int32_t *) node->op_params;
    const int n_dims = prm[1], mode = prm[2], n_ctx_orig = prm[4];
    float freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow;
    memcpy(&freq_base, prm + 5, 4); memcpy(&freq_scale, prm + 6, 4); memcpy(&ext_factor, prm + 7, 4);
    memcpy(&attn_factor, prm + 8, 4); memcpy(&beta_fast, prm + 9, 4); memcpy(&beta_slow, prm + 10, 4);
    // the two host-side derivations of ggml_compute_forward_rope_f32 (ops.cpp:5584-5587), with the reference's own libm / helper
    const float theta_scale = powf(freq_base, -2.0f / n_dims);
    float corr_dims[2];
    ggml_rope_yarn_corr_dims(n_dims, n_ctx_orig, freq_base, beta_fast, beta_slow, corr_dims);
    const int64_t x_nb[3] = {(int64_t) x->nb[1], (int64_t) x->nb[2], (int64_t) x->nb[3]};
    const int64_t d_nb[3] = {(int64_t) node->nb[1], (int64_t) node->nb[2], (int64_t) node->nb[3]};
    const int rc = q4k_op_rope((const float *) x->data, x_nb, (const int32_t *) pos->data, ff ? (const float *) ff->data : nullptr, (float *) node->data,
                               d_nb, x->ne, n_dims, mode == GGML_ROPE_TYPE_NEOX, theta_scale, freq_scale, ext_factor, attn_factor, corr_dims[0],
                               corr_dims[1], ctx->stream);
    return rc == Q4K_OK ? GGML_STATUS_SUCCESS : op_failed("q4k_op_rope", node, rc);
}

bool is_noop(const ggml_tensor * node) {
    return node->op == GGML_OP_NONE || node->op == GGML_OP_RESHAPE || node->op == GGML_OP_VIEW || node->op == GGML_OP_PERMUTE ||
           node->op == GGML_OP_TRANSPOSE;
}

// Q8_K scratch for the largest mul_mat of the graph, sized BEFORE any launch (no allocation, no synchronize between launches:
// the launch sequence may be under stream capture).  Grow-only; growing invalidates captured launches (they hold the pointer).
ggml_status ensure_workspace(backend_context * ctx, const ggml_cgraph * cgraph) {
    size_t need = 0;
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        const ggml_tensor * node = cgraph->nodes[i];
        if (node->op != GGML_OP_MUL_MAT || !node->src[0] || !node->src[1]) continue;
        const int64_t K = node->src[0]->ne[0], N = node->src[1]->ne[1];
        const size_t n = node->src[0]->type == GGML_TYPE_Q6_K ? q6k_mul_mat_workspace_bytes(K, N) : q4k_mul_mat_workspace_bytes(K, N);
        if (n > need) need = n;
    }
    if (need <= ctx->workspace_bytes) return GGML_STATUS_SUCCESS;
    // in-flight kernels may still read the old one
    if (q4k_stream_synchronize(ctx->dev->cuda_device, ctx->stream) != Q4K_OK) return GGML_STATUS_FAILED;
    q4k_mem_free(ctx->dev->cuda_device, ctx->workspace);
    ctx->workspace = q4k_mem_alloc(ctx->dev->cuda_device, need);
    ctx->workspace_bytes = ctx->workspace ? need : 0;
    ++ctx->ws_generation;
    if (!ctx->workspace) {
        GGML_LOG_ERROR("Q4K_B200: cannot allocate %zu bytes of Q8_K workspace\n", need);
        return GGML_STATUS_ALLOC_FAILED;
    }
    return GGML_STATUS_SUCCESS;
}

// What a captured launch sequence depends on: per launching node the op, types, device pointers, shapes and strides.
void graph_signature(const ggml_cgraph * cgraph, std::vector<uint64_t> & sig, int & launching) {
    sig.clear();
    launching = 0;
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        const ggml_tensor * node = cgraph->nodes[i];
        if (is_noop(node)) continue;
        ++launching;
        const ggml_tensor * t[4] = {node, node->src[0], node->src[1], node->src[2]};
        sig.push_back((uint64_t) node->op | ((uint64_t) (uint32_t) node->op_params[0] << 32));      // op_params[0]: eps of RMS_NORM, the GLU kind
        if (node->op == GGML_OP_ROPE) for (int k = 1; k < 11; ++k) sig.push_back((uint64_t) (uint32_t) node->op_params[k]);
        for (const ggml_tensor * x : t) {
            if (!x) { sig.push_back(0); continue; }
            sig.push_back((uint64_t) (uintptr_t) x->data);
            sig.push_back((uint64_t) x->type);
            for (int d = 0; d < GGML_MAX_DIMS; ++d) { sig.push_back((uint64_t) x->ne[d]); sig.push_back((uint64_t) x->nb[d]); }
        }
        // consecutive prefill mul_mats share the quantized activations only if they name the SAME src1 tensor
        sig.push_back(i > 0 && cgraph->nodes[i - 1]->src[1] == node->src[1] ? 1u : 0u);
    }
}

ggml_status enqueue_nodes(backend_context * ctx, ggml_cgraph * cgraph) {
    const ggml_tensor * last_quantized = nullptr;      // src1 whose prefill-quantized copy is in the workspace right now
    const void * last_quantized_data = nullptr;
    std::vector<char> done(cgraph->n_nodes, 0);         // mul_mats that already went out with an earlier node's group launch
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        ggml_tensor * node = cgraph->nodes[i];
        if (done[i]) continue;
        switch (node->op) {
            case GGML_OP_MUL_MAT: {
                if (!mul_mat_supported(node)) {
                    GGML_ABORT("Q4K_B200: %s: MUL_MAT node outside the supports_op contract reached graph_compute", node->name);
                }
                const int fused = try_fused_group(ctx, cgraph, i, done);
                if (fused < 0) return GGML_STATUS_FAILED;
                if (fused > 0) { last_quantized = nullptr; break; }
                // consecutive prefill mul_mats on one src1 (q/k/v, gate/up) share the quantized copy of the activations
                const ggml_tensor * s0 = node->src[0], * s1 = node->src[1];
                const bool plain2d = s0->ne[2] == 1 && s0->ne[3] == 1 && s1->ne[2] == 1 && s1->ne[3] == 1 && !tensor_is_split(s0);
                const bool reuse = plain2d && s0->type == GGML_TYPE_Q4_K && last_quantized == s1 && last_quantized_data == s1->data;
                const ggml_status st = compute_mul_mat(ctx, node, reuse);
                if (st != GGML_STATUS_SUCCESS) return st;
                if (plain2d && s0->type == GGML_TYPE_Q4_K) { last_quantized = s1; last_quantized_data = s1->data; }
                else { last_quantized = nullptr; last_quantized_data = nullptr; }
            } break;
            case GGML_OP_ADD:
            case GGML_OP_RMS_NORM: {
                const ggml_tensor * add, * norm, * mul;
                const int n = norm_chain(cgraph, i, add, norm, mul);       // [ADD] RMS_NORM [MUL] in one launch
                last_quantized = nullptr;
                if (n > 0) {
                    const ggml_status st = compute_rms_norm(ctx, add, norm, mul);
                    if (st != GGML_STATUS_SUCCESS) return st;
                    i += n - 1;
                    break;
                }
                if (node->op == GGML_OP_RMS_NORM) GGML_ABORT("Q4K_B200: %s: RMS_NORM node outside the supports_op contract", node->name);
            }
            [[fallthrough]];      // a plain ADD
            case GGML_OP_MUL: {
                if (!binary_supported(node)) GGML_ABORT("Q4K_B200: %s: %s node outside the supports_op contract", node->name, ggml_op_name(node->op));
                last_quantized = nullptr;
                const ggml_status st = compute_binary(ctx, node);
                if (st != GGML_STATUS_SUCCESS) return st;
            } break;
            case GGML_OP_GLU: {
                if (!glu_supported(node)) GGML_ABORT("Q4K_B200: %s: GLU node outside the supports_op contract", node->name);
                last_quantized = nullptr;
                const ggml_status st = compute_glu(ctx, node);
                if (st != GGML_STATUS_SUCCESS) return st;
            } break;
            case GGML_OP_ROPE: {
                if (!rope_supported(node)) GGML_ABORT("Q4K_B200: %s: ROPE node outside the supports_op contract", node->name);
                last_quantized = nullptr;
                const ggml_status st = compute_rope(ctx, node);
                if (st != GGML_STATUS_SUCCESS) return st;
            } break;
            case GGML_OP_GET_ROWS: {
                if (!get_rows_supported(node)) GGML_ABORT("Q4K_B200: %s: GET_ROWS node outside the supports_op contract", node->name);
                last_quantized = nullptr;
                const ggml_status st = compute_get_rows(ctx, node);
                if (st != GGML_STATUS_SUCCESS) return st;
            } break;
            case GGML_OP_NONE:
            case GGML_OP_RESHAPE:
            case GGML_OP_VIEW:
            case GGML_OP_PERMUTE:
            case GGML_OP_TRANSPOSE:
                break;
            default:
                GGML_ABORT("Q4K_B200: unsupported op %s reached graph_compute", ggml_op_desc(node));
        }
    }
    return GGML_STATUS_SUCCESS;
}

// Runs the graph through `cache`: plain launches the first time a launch sequence is seen, captured into a CUDA graph the
// second time (78 dependent launches of a decode token cost one graph launch after that), replayed from then on.
ggml_status compute_cached(backend_context * ctx, ggml_cgraph * cgraph, graph_cache & cache) {
    if (q4k_set_device(ctx->dev->cuda_device) != Q4K_OK) return GGML_STATUS_FAILED;
    const ggml_status ws = ensure_workspace(ctx, cgraph);
    if (ws != GGML_STATUS_SUCCESS) return ws;
    static thread_local std::vector<uint64_t> sig;
    int launching = 0;
    graph_signature(cgraph, sig, launching);
    if (launching < 2) return enqueue_nodes(ctx, cgraph);                    // nothing to amortise
    for (int i = 0; i < cgraph->n_nodes; ++i)                                // shard launches fork onto other devices' streams with events:
        if (cgraph->nodes[i]->op == GGML_OP_MUL_MAT && tensor_is_split(cgraph->nodes[i]->src[0])) return enqueue_nodes(ctx, cgraph);   // plain launches
    const bool same = sig == cache.sig;
    if (same && cache.graph && cache.ws_generation == ctx->ws_generation) {
        if (q4k_graph_launch(cache.graph, ctx->stream) == Q4K_OK) return GGML_STATUS_SUCCESS;
        GGML_LOG_ERROR("Q4K_B200: graph launch failed (cudaError %d)\n", q4k_last_cuda_error());
        return GGML_STATUS_FAILED;
    }
    if (!same) {
        q4k_graph_destroy(cache.graph);
        cache.graph = nullptr; cache.sig = sig; cache.seen = 0; cache.failed = false;
    } else if (cache.graph) {                                                // same launches, stale workspace pointer
        q4k_graph_destroy(cache.graph);
        cache.graph = nullptr;
    }
    if (++cache.seen < 2 || cache.failed) return enqueue_nodes(ctx, cgraph);
    if (q4k_graph_begin_capture(ctx->stream) != Q4K_OK) { cache.failed = true; return enqueue_nodes(ctx, cgraph); }
    const ggml_status st = enqueue_nodes(ctx, cgraph);
    q4k_graph * g = nullptr;
    const int rc = q4k_graph_end_capture(ctx->stream, st == GGML_STATUS_SUCCESS, &g);
    if (st != GGML_STATUS_SUCCESS || rc != Q4K_OK || !g) {
        cache.failed = true;                                                 // e.g. a launch that cannot be captured: plain launches
        return st != GGML_STATUS_SUCCESS ? st : enqueue_nodes(ctx, cgraph);
    }
    cache.graph = g;
    cache.ws_generation = ctx->ws_generation;
    return q4k_graph_launch(g, ctx->stream) == Q4K_OK ? GGML_STATUS_SUCCESS : GGML_STATUS_FAILED;
}

ggml_status backend_graph_compute(ggml_backend_t backend, ggml_cgraph * cgraph) {
    auto * ctx = (backend_context *) backend->context;
    return compute_cached(ctx, cgraph, ctx->cache);      // asynchronous: results are complete after synchronize()
}

// backend_i.graph_plan_* (L/ggml/src/ggml-backend-impl.h:100-108; the CPU backend's: L/ggml/src/ggml-cpu/ggml-cpu.cpp:128-165)
ggml_backend_graph_plan_t backend_graph_plan_create(ggml_backend_t, const ggml_cgraph * cgraph) {
    auto * plan = new graph_plan;
    plan->nodes.assign(cgraph->nodes, cgraph->nodes + cgraph->n_nodes);
    plan->cgraph = *cgraph;
    plan->cgraph.nodes = plan->nodes.data();
    return plan;
}

void backend_graph_plan_free(ggml_backend_t backend, ggml_backend_graph_plan_t plan) {
    auto * ctx = (backend_context *) backend->context;
    q4k_stream_synchronize(ctx->dev->cuda_device, ctx->stream);
    delete (graph_plan *) plan;
}

void backend_graph_plan_update(ggml_backend_t, ggml_backend_graph_plan_t plan_, const ggml_cgraph * cgraph) {
    auto * plan = (graph_plan *) plan_;
    plan->nodes.assign(cgraph->nodes, cgraph->nodes + cgraph->n_nodes);
    plan->cgraph = *cgraph;
    plan->cgraph.nodes = plan->nodes.data();              // the cache notices by itself if the launches changed
}

ggml_status backend_graph_plan_compute(ggml_backend_t backend, ggml_backend_graph_plan_t plan_) {
    auto * ctx = (backend_context *) backend->context;
    auto * plan = (graph_plan *) plan_;
    // first compute: plain launches (kernel modules load, per-kernel attributes are set); second: captured; then replayed
    return compute_cached(ctx, &plan->cgraph, plan->cache);
}

const ggml_backend_i backend_iface = {
    /* .get_name           = */ backend_get_name,
    /* .free               = */ backend_free,
    /* .set_tensor_async   = */ backend_set_tensor_async,
    /* .get_tensor_async   = */ backend_get_tensor_async,
    /* .cpy_tensor_async   = */ nullptr,
    /* .synchronize        = */ backend_synchronize,
    /* .graph_plan_create  = */ backend_graph_plan_create,
    /* .graph_plan_free    = */ backend_graph_plan_free,
    /* .graph_plan_update  = */ backend_graph_plan_update,
    /* .graph_plan_compute = */ backend_graph_plan_compute,
    /* .graph_compute      = */ backend_graph_compute,
    /* .event_record       = */ nullptr,
    /* .event_wait         = */ nullptr,
    /* .graph_optimize     = */ nullptr,
};

ggml_guid_t backend_guid() {
    static ggml_guid guid = {0x51, 0x34, 0x4b, 0x5f, 0x42, 0x32, 0x30, 0x30, 0x9a, 0x10, 0x0a, 0x73, 0x6d, 0x31, 0x30, 0x30};
    return &guid;
}

}  // namespace

namespace {

// ------------------------------------------------------------------------------------------ device
const char * device_get_name(ggml_backend_dev_t dev) { return ((device_context *) dev->context)->name.c_str(); }
const char * device_get_description(ggml_backend_dev_t dev) { return ((device_context *) dev->context)->description.c_str(); }

void device_get_memory(ggml_backend_dev_t dev, size_t * free, size_t * total) {
    auto * ctx = (device_context *) dev->context;
    size_t f = 0, t = 0;
    q4k_device_info(ctx->cuda_device, nullptr, 0, &f, &t, nullptr);
    *free = f;
    *total = t;
}

enum ggml_backend_dev_type device_get_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }

ggml_backend_buffer_type_t device_get_buffer_type(ggml_backend_dev_t dev) {
    return &((device_context *) dev->context)->buft;
}

ggml_backend_buffer_type_t device_get_host_buffer_type(ggml_backend_dev_t dev) {
    return &((device_context *) dev->context)->host_buft;
}

void device_get_props(ggml_backend_dev_t dev, ggml_backend_dev_props * props) {
    props->name        = device_get_name(dev);
    props->description = device_get_description(dev);
    props->type        = device_get_type(dev);
    device_get_memory(dev, &props->memory_free, &props->memory_total);
    props->caps = {
        /* .async                = */ true,
        /* .host_buffer          = */ true,
        /* .buffer_from_host_ptr = */ false,
        /* .events               = */ false,
    };
}

ggml_backend_t device_init_backend(ggml_backend_dev_t dev, const char *) {
    auto * dctx = (device_context *) dev->context;
    void * stream = q4k_stream_create(dctx->cuda_device);
    if (!stream) {
        GGML_LOG_ERROR("Q4K_B200: cannot create a stream on device %d (cudaError %d)\n", dctx->cuda_device, q4k_last_cuda_error());
        return nullptr;
    }
    auto * ctx = new backend_context;
    ctx->dev = dctx;
    ctx->stream = stream;
    return new ggml_backend{
        /* .guid    = */ backend_guid(),
        /* .iface   = */ backend_iface,
        /* .device  = */ dev,
        /* .context = */ ctx,
    };
}

bool cheap_ops_enabled() {
    static const bool on = [] { const char * e = getenv("Q4K_B200_CHEAP_OPS"); return !(e && e[0] == '0'); }();
    return on;
}

bool device_supports_op(ggml_backend_dev_t, const ggml_tensor * op) {
    if (op->op != GGML_OP_MUL_MAT)       // a split tensor has no address: only MUL_MAT (as src0) knows how to read it
        for (int k = 0; k < GGML_MAX_SRC; ++k) if (tensor_is_split(op->src[k])) return false;
    if (op->op == GGML_OP_MUL_MAT && tensor_is_split(op->src[1])) return false;
    switch (op->op) {
        // test-backend-ops asks about every tensor of the context, leaves and views included
        // (L/tests/test-backend-ops.cpp:1152-1161); these are no-ops for a backend
        case GGML_OP_NONE:
        case GGML_OP_RESHAPE:
        case GGML_OP_VIEW:
        case GGML_OP_PERMUTE:
        case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_MUL_MAT:
            return mul_mat_supported(op);
        // the cheap ops of a decoder layer (Q4K_B200_CHEAP_OPS=0 leaves them to the CPU backend)
        case GGML_OP_RMS_NORM:
            return cheap_ops_enabled() && rms_norm_supported(op);
        case GGML_OP_ADD:
        case GGML_OP_MUL:
            return cheap_ops_enabled() && binary_supported(op);
        case GGML_OP_GLU:
            return cheap_ops_enabled() && glu_supported(op);
        case GGML_OP_GET_ROWS:
            return cheap_ops_enabled() && get_rows_supported(op);
        case GGML_OP_ROPE:
            return cheap_ops_enabled() && rope_supported(op);
        default:
            return false;
    }
}

bool device_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    if (buft_is_split(buft)) return true;              // any of our devices can be the main device of a split matrix
    return buft_is_ours(buft) && buft->context == dev->context;
}

const ggml_backend_device_i device_iface = {
    /* .get_name             = */ device_get_name,
    /* .get_description      = */ device_get_description,
    /* .get_memory           = */ device_get_memory,
    /* .get_type             = */ device_get_type,
    /* .get_props            = */ device_get_props,
    /* .init_backend         = */ device_init_backend,
    /* .get_buffer_type      = */ device_get_buffer_type,
    /* .get_host_buffer_type = */ device_get_host_buffer_type,
    /* .buffer_from_host_ptr = */ nullptr,
    /* .supports_op          = */ device_supports_op,
    /* .supports_buft        = */ device_supports_buft,
    /* .offload_op           = */ nullptr,
    /* .event_new            = */ nullptr,
    /* .event_free           = */ nullptr,
    /* .event_synchronize    = */ nullptr,
};

// ------------------------------------------------------------------------------------------ registry
const char * reg_get_name(ggml_backend_reg_t) { return "Q4K_B200"; }
size_t reg_get_device_count(ggml_backend_reg_t) { return devices().size(); }
ggml_backend_dev_t reg_get_device(ggml_backend_reg_t, size_t index) {
    GGML_ASSERT(index < devices().size());
    return &devices()[index];
}
void * reg_get_proc_address(ggml_backend_reg_t, const char * name) {
    if (strcmp(name, "ggml_backend_split_buffer_type") == 0) return (void *) split_buffer_type;
    return nullptr;
}

const ggml_backend_reg_i reg_iface = {
    /* .get_name         = */ reg_get_name,
    /* .get_device_count = */ reg_get_device_count,
    /* .get_device       = */ reg_get_device,
    /* .get_proc_address = */ reg_get_proc_address,
};

int count_sm100_devices() {
    int n = 0;
    const int total = q4k_device_count();
    for (int d = 0; d < total; ++d) n += q4k_device_ok(d) ? 1 : 0;
    return n;
}

ggml_backend_reg_t q4k_b200_reg() {
    static ggml_backend_reg reg = {
        /* .api_version = */ GGML_BACKEND_API_VERSION,
        /* .iface       = */ reg_iface,
        /* .context     = */ nullptr,
    };
    static std::once_flag once;
    std::call_once(once, [] {
        const int total = q4k_device_count();
        auto & dctx = device_contexts();
        dctx.reserve(total);                    // pointers into these vectors are handed out: never reallocate
        for (int d = 0; d < total; ++d) {
            if (!q4k_device_ok(d)) continue;    // sm_100 only: the kernels are compiled for sm_100a and nothing else
            char name[256] = "";
            int sms = 0;
            q4k_device_info(d, name, sizeof name, nullptr, nullptr, &sms);
            device_context c;
            c.cuda_device = d;
            c.name = "Q4K_B200_" + std::to_string(dctx.size());
            c.description = std::string(name) + " (" + std::to_string(sms) + " SMs, " + q4k_build_info() + ")";
            dctx.push_back(c);
        }
        auto & devs = devices();
        devs.reserve(dctx.size());
        for (auto & c : dctx) {
            devs.push_back(ggml_backend_device{device_iface, &reg, &c});
            c.buft = ggml_backend_buffer_type{buft_iface, &devs.back(), &c};
            c.host_buft = ggml_backend_buffer_type{host_buft_iface, &devs.back(), &c};
        }
        GGML_LOG_INFO("Q4K_B200: %zu sm_100 device(s); claims MUL_MAT(Q4_K | Q6_K, F32)%s\n", devs.size(),
                      cheap_ops_enabled() ? " and the F32 ops around it (RMS_NORM, ADD, MUL, GLU/SWIGLU, GET_ROWS, ROPE)" : " only");
    });
    return &reg;
}

}  // namespace

// DL entry points (L/ggml/src/ggml-backend-impl.h:214-251); declared here without GGML_BACKEND_DL so the file
// builds the same way in or out of the reference tree
extern "C" {
__attribute__((visibility("default"))) ggml_backend_reg_t ggml_backend_init(void) { return q4k_b200_reg(); }
__attribute__((visibility("default"))) int ggml_backend_score(void) { return count_sm100_devices() > 0 ? 100 : 0; }
}
