// c_api_mem.cu — device memory, pinned host memory and stream entries of include/q4k_b200.h.
// Thin pass-throughs to the CUDA runtime so that host code above the C ABI (the ggml backend plugin) needs no
// CUDA headers.  Errors map to Q4K_ERR_CUDA with the code kept for q4k_last_cuda_error().
#include "../../include/q4k_b200.h"

#include <cstdio>
#include <cstring>

#include "q4k_kernels.h"

namespace q4k {
int note_cuda_error(cudaError_t e);      // c_api.cu
}
using q4k::note_cuda_error;

#define Q4K_TRY(expr)                                              \
    do {                                                           \
        const cudaError_t e_ = (expr);                             \
        if (e_ != cudaSuccess) return note_cuda_error(e_);         \
    } while (0)

extern "C" {

int q4k_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int q4k_set_device(int device) {
    Q4K_TRY(cudaSetDevice(device));
    return Q4K_OK;
}

int q4k_device_info(int device, char *name, size_t name_len, size_t *free_bytes, size_t *total_bytes, int *sm_count) {
    cudaDeviceProp prop;
    Q4K_TRY(cudaGetDeviceProperties(&prop, device));
    if (name && name_len) snprintf(name, name_len, "%s", prop.name);
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (free_bytes || total_bytes) {
        int cur = 0;
        Q4K_TRY(cudaGetDevice(&cur));
        Q4K_TRY(cudaSetDevice(device));
        size_t f = 0, t = 0;
        const cudaError_t e = cudaMemGetInfo(&f, &t);
        cudaSetDevice(cur);
        if (e != cudaSuccess) return note_cuda_error(e);
        if (free_bytes) *free_bytes = f;
        if (total_bytes) *total_bytes = t;
    }
    return Q4K_OK;
}

void *q4k_mem_alloc(int device, size_t bytes) {
    void *p = nullptr;
    if (cudaSetDevice(device) != cudaSuccess) return nullptr;
    const cudaError_t e = cudaMalloc(&p, bytes ? bytes : 1);
    if (e != cudaSuccess) { note_cuda_error(e); (void)cudaGetLastError(); return nullptr; }
    return p;
}

int q4k_mem_free(int device, void *ptr_dev) {
    if (!ptr_dev) return Q4K_OK;
    Q4K_TRY(cudaSetDevice(device));
    Q4K_TRY(cudaFree(ptr_dev));
    return Q4K_OK;
}

int q4k_mem_set(int device, void *dst_dev, int value, size_t bytes, void *stream) {
    if (bytes == 0) return Q4K_OK;
    if (!dst_dev) return Q4K_ERR_BAD_POINTER;
    Q4K_TRY(cudaSetDevice(device));
    if (stream) Q4K_TRY(cudaMemsetAsync(dst_dev, value, bytes, (cudaStream_t)stream));
    else { Q4K_TRY(cudaMemset(dst_dev, value, bytes)); Q4K_TRY(cudaDeviceSynchronize()); }
    return Q4K_OK;
}

static int copy(int device, void *dst, const void *src, size_t bytes, cudaMemcpyKind kind, void *stream) {
    if (bytes == 0) return Q4K_OK;
    if (!dst || !src) return Q4K_ERR_BAD_POINTER;
    Q4K_TRY(cudaSetDevice(device));
    if (stream) Q4K_TRY(cudaMemcpyAsync(dst, src, bytes, kind, (cudaStream_t)stream));
    else Q4K_TRY(cudaMemcpy(dst, src, bytes, kind));
    return Q4K_OK;
}

int q4k_mem_h2d(int device, void *dst_dev, const void *src_host, size_t bytes, void *stream) {
    return copy(device, dst_dev, src_host, bytes, cudaMemcpyHostToDevice, stream);
}
int q4k_mem_d2h(int device, void *dst_host, const void *src_dev, size_t bytes, void *stream) {
    return copy(device, dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, stream);
}
int q4k_mem_d2d(int device, void *dst_dev, const void *src_dev, size_t bytes, void *stream) {
    return copy(device, dst_dev, src_dev, bytes, cudaMemcpyDeviceToDevice, stream);
}

void *q4k_host_alloc_pinned(size_t bytes) {
    void *p = nullptr;
    const cudaError_t e = cudaMallocHost(&p, bytes ? bytes : 1);
    if (e != cudaSuccess) { note_cuda_error(e); (void)cudaGetLastError(); return nullptr; }
    return p;
}

int q4k_host_free_pinned(void *ptr_host) {
    if (!ptr_host) return Q4K_OK;
    Q4K_TRY(cudaFreeHost(ptr_host));
    return Q4K_OK;
}

void *q4k_stream_create(int device) {
    cudaStream_t s = nullptr;
    if (cudaSetDevice(device) != cudaSuccess) return nullptr;
    // blocking flag on purpose: synchronous copies on the legacy default stream then order against this stream
    const cudaError_t e = cudaStreamCreate(&s);
    if (e != cudaSuccess) { note_cuda_error(e); return nullptr; }
    return (void *)s;
}

int q4k_stream_destroy(int device, void *stream) {
    if (!stream) return Q4K_OK;
    Q4K_TRY(cudaSetDevice(device));
    Q4K_TRY(cudaStreamDestroy((cudaStream_t)stream));
    return Q4K_OK;
}

int q4k_stream_synchronize(int device, void *stream) {
    Q4K_TRY(cudaSetDevice(device));
    Q4K_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    return Q4K_OK;
}

// ---- captured launch sequences (backend_i.graph_plan_* and the plugin's cached graph_compute) ------------------------
struct q4k_graph {
    cudaGraphExec_t exec;
    int64_t launches;       // kernels of this library inside the graph (for q4k_kernel_launches)
};
static thread_local int64_t t_capture_launches0 = 0;

int q4k_graph_begin_capture(void *stream) {
    if (!stream) return Q4K_ERR_BAD_POINTER;           // the legacy default stream cannot be captured
    t_capture_launches0 = q4k::launch_count();
    Q4K_TRY(cudaStreamBeginCapture((cudaStream_t)stream, cudaStreamCaptureModeThreadLocal));
    return Q4K_OK;
}

int q4k_graph_end_capture(void *stream, int keep, q4k_graph **graph_out) {
    cudaGraph_t g = nullptr;
    const cudaError_t e = cudaStreamEndCapture((cudaStream_t)stream, &g);
    if (e != cudaSuccess) { (void)cudaGetLastError(); if (g) cudaGraphDestroy(g); return note_cuda_error(e); }
    if (!keep || !graph_out) { if (g) cudaGraphDestroy(g); return Q4K_OK; }      // the caller abandons the capture (a launch failed)
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ei = cudaGraphInstantiate(&exec, g, 0);
    cudaGraphDestroy(g);
    if (ei != cudaSuccess) { (void)cudaGetLastError(); return note_cuda_error(ei); }
    *graph_out = new q4k_graph{exec, q4k::launch_count() - t_capture_launches0};
    return Q4K_OK;
}

int q4k_graph_launch(q4k_graph *graph, void *stream) {
    if (!graph) return Q4K_ERR_BAD_POINTER;
    Q4K_TRY(cudaGraphLaunch(graph->exec, (cudaStream_t)stream));
    q4k::add_launches(graph->launches);
    return Q4K_OK;
}

void q4k_graph_destroy(q4k_graph *graph) {
    if (!graph) return;
    cudaGraphExecDestroy(graph->exec);
    delete graph;
}

}  // extern "C"
