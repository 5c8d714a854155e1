// split.cu — row-sharded weight matrices over the GPUs of ONE process (q4k_split_* of include/q4k_b200.h).
// The multi-GPU path of SURVEY.md §8e inside the C / C++ boundary: what ggml-cuda's split buffer does for --split-mode row
// (L/ggml/src/ggml-cuda/ggml-cuda.cu:774-787 get_row_split, :795-1000 the split buffer, :1672 / :1720 src1 / dst peer copies),
// re-cut for this path: rows of a weight matrix are dealt to the devices in contiguous, tile-aligned ranges at upload time
// (q4k_split_set = buffer_i.set_tensor); a mul_mat launches each device's shard on that device's own stream, the shard kernels
// READ the activations from the main device's memory and WRITE their result rows straight into the main device's dst over
// NVLink peer mappings (no staging copies: the decode kernel's epilogue is the gather), and the main stream joins them with
// events.  Results are the single-GPU results bit for bit: a row's bits do not depend on where it sits (gemv_q4k_mma.cu).
//
// The one-process-per-GPU path with the LL exchange (q4k_mul_mat_peers_ll; q4k_b200/shard.py, bench.py --gpus N) stays the fast
// path for the benchmark — it all-gathers to EVERY rank inside the launch; this one gathers to ONE device, which is what a ggml
// graph wants (dst lives in the main backend's compute buffer).
#include "../../include/q4k_b200.h"

#include <cstdint>
#include <cstring>
#include <mutex>
#include <vector>

#include "q4k_common.cuh"
#include "q4k_kernels.h"

namespace q4k {
int note_cuda_error(cudaError_t e);
}
using q4k::note_cuda_error;

#define Q4K_TRY(expr)                                              \
    do {                                                           \
        const cudaError_t e_ = (expr);                             \
        if (e_ != cudaSuccess) return note_cuda_error(e_);         \
    } while (0)

struct q4k_split {
    struct Part {
        int device;
        int64_t row_low, row_high;
        void *w = nullptr;               // rows [row_low, row_high) of the matrix, on `device`
    };
    std::vector<Part> parts;
    int64_t M, row_bytes;
};

namespace {

// Per device, shared by every split matrix of the process: the stream the device's shards run on, the event that marks a
// shard done, the Q8_K workspace (grown on demand), and — for a device in the main role — the event recorded on the caller's
// stream when the activations are ready.  One mul_mat at a time per main stream is in flight on them (stream order).
struct DevCtx {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr, ready = nullptr;
    void *ws = nullptr;
    size_t ws_bytes = 0;
};
std::mutex g_dev_mutex;
DevCtx g_dev[64];

DevCtx *dev_ctx(int device) {             // current device must be `device`
    if (device < 0 || device >= 64) return nullptr;
    std::lock_guard<std::mutex> lock(g_dev_mutex);
    DevCtx &d = g_dev[device];
    if (!d.stream) {
        if (cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&d.done, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&d.ready, cudaEventDisableTiming) != cudaSuccess) {
            (void)cudaGetLastError();
            return nullptr;
        }
    }
    return &d;
}

int enable_peer(int from, int to) {       // kernels on `from` may dereference pointers into `to`'s memory
    if (from == to) return Q4K_OK;
    int can = 0;
    Q4K_TRY(cudaDeviceCanAccessPeer(&can, from, to));
    if (!can) return Q4K_ERR_UNSUPPORTED;
    Q4K_TRY(cudaSetDevice(from));
    const cudaError_t e = cudaDeviceEnablePeerAccess(to, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); return Q4K_OK; }
    return e == cudaSuccess ? Q4K_OK : note_cuda_error(e);
}

}  // namespace

extern "C" {

q4k_split *q4k_split_create(const int *devices, int n_devices, const float *fractions, int64_t M, int64_t row_bytes, int row_granule) {
    if (!devices || n_devices < 1 || n_devices > 16 || M < 0 || row_bytes <= 0 || row_granule < 1) return nullptr;
    int cur = 0;
    if (cudaGetDevice(&cur) != cudaSuccess) return nullptr;
    auto *s = new q4k_split;
    s->M = M; s->row_bytes = row_bytes;
    // cumulative starts: fractions[i] = share of device i (NULL or all zero = equal shares), rounded down to the granule
    double total = 0;
    for (int i = 0; i < n_devices; ++i) total += fractions ? (double)fractions[i] : 0.0;
    double acc = 0;
    bool ok = true;
    for (int i = 0; i < n_devices && ok; ++i) {
        q4k_split::Part p;
        p.device = devices[i];
        const double lo = total > 0 ? acc / total : (double)i / n_devices;
        acc += fractions ? (double)fractions[i] : 0.0;
        const double hi = total > 0 ? acc / total : (double)(i + 1) / n_devices;
        p.row_low = i == 0 ? 0 : (int64_t)(lo * (double)M) / row_granule * row_granule;
        p.row_high = i == n_devices - 1 ? M : (int64_t)(hi * (double)M) / row_granule * row_granule;
        if (p.row_high < p.row_low) p.row_high = p.row_low;
        ok = cudaSetDevice(p.device) == cudaSuccess;
        const size_t bytes = (size_t)(p.row_high - p.row_low) * (size_t)row_bytes;
        if (ok && bytes) ok = cudaMalloc(&p.w, bytes) == cudaSuccess;
        s->parts.push_back(p);
    }
    if (!ok) { (void)cudaGetLastError(); q4k_split_destroy(s); s = nullptr; }
    cudaSetDevice(cur);
    return s;
}

void q4k_split_destroy(q4k_split *s) {
    if (!s) return;
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto &p : s->parts) {
        if (cudaSetDevice(p.device) != cudaSuccess) continue;
        if (DevCtx *d = dev_ctx(p.device)) cudaStreamSynchronize(d->stream);      // a shard launch may still read the weights
        if (p.w) cudaFree(p.w);
    }
    cudaSetDevice(cur);
    delete s;
}

int q4k_split_parts(const q4k_split *s) { return s ? (int)s->parts.size() : 0; }

int q4k_split_range(const q4k_split *s, int part, int *device, int64_t *row_low, int64_t *row_high) {
    if (!s || part < 0 || part >= (int)s->parts.size()) return Q4K_ERR_BAD_SHAPE;
    if (device) *device = s->parts[part].device;
    if (row_low) *row_low = s->parts[part].row_low;
    if (row_high) *row_high = s->parts[part].row_high;
    return Q4K_OK;
}

// bytes [offset, offset + size) of the full row-major matrix <-> the shards (synchronous, like buffer_i.set_tensor / get_tensor)
static int split_copy(q4k_split *s, void *host, size_t offset, size_t size, bool to_device) {
    if (!s || (!host && size)) return Q4K_ERR_BAD_POINTER;
    if (offset + size > (size_t)s->M * (size_t)s->row_bytes) return Q4K_ERR_BAD_SHAPE;
    int cur = 0;
    Q4K_TRY(cudaGetDevice(&cur));
    int rc = Q4K_OK;
    for (auto &p : s->parts) {
        const size_t lo = (size_t)p.row_low * (size_t)s->row_bytes, hi = (size_t)p.row_high * (size_t)s->row_bytes;
        const size_t a = offset > lo ? offset : lo, b = offset + size < hi ? offset + size : hi;
        if (a >= b) continue;
        cudaError_t e = cudaSetDevice(p.device);
        if (e == cudaSuccess)
            e = to_device ? cudaMemcpy((char *)p.w + (a - lo), (const char *)host + (a - offset), b - a, cudaMemcpyHostToDevice)
                          : cudaMemcpy((char *)host + (a - offset), (const char *)p.w + (a - lo), b - a, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { rc = note_cuda_error(e); break; }
    }
    cudaSetDevice(cur);
    return rc;
}

int q4k_split_set(q4k_split *s, const void *host, size_t offset, size_t size) { return split_copy(s, const_cast<void *>(host), offset, size, true); }
int q4k_split_get(q4k_split *s, void *host, size_t offset, size_t size) { return split_copy(s, host, offset, size, false); }

int q4k_split_mul_mat(q4k_split *s, int q6k, int64_t K, const float *x_dev, int64_t N, int64_t x_row_stride_bytes, float *dst_dev,
                      int64_t dst_col_stride_bytes, int main_device, void *main_stream) {
    if (!s || K <= 0 || K % Q4K_QK_K != 0 || N < 0) return Q4K_ERR_BAD_SHAPE;
    if (s->M == 0 || N == 0) return Q4K_OK;
    if (!x_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
    const int64_t blk = q6k ? Q4K_BLOCK_Q6K : Q4K_BLOCK_Q4K;
    if (s->row_bytes != K / Q4K_QK_K * blk || dst_col_stride_bytes < s->M * 4) return Q4K_ERR_BAD_SHAPE;
    int cur = 0;
    Q4K_TRY(cudaGetDevice(&cur));
    // the activations are ready when everything queued on the caller's stream so far has run
    Q4K_TRY(cudaSetDevice(main_device));
    DevCtx *md = dev_ctx(main_device);
    if (!md) return Q4K_ERR_CUDA;
    Q4K_TRY(cudaEventRecord(md->ready, (cudaStream_t)main_stream));
    int rc = Q4K_OK;
    const size_t need = q6k ? q6k_mul_mat_workspace_bytes(K, N) : q4k_mul_mat_workspace_bytes(K, N);
    for (auto &p : s->parts) {
        const int64_t rows = p.row_high - p.row_low;
        if (rows == 0) continue;
        if ((rc = enable_peer(p.device, main_device)) != Q4K_OK) break;
        cudaError_t e = cudaSetDevice(p.device);
        DevCtx *d = e == cudaSuccess ? dev_ctx(p.device) : nullptr;
        if (!d) { rc = Q4K_ERR_CUDA; break; }
        if (need > d->ws_bytes) {                               // grown on demand (synchronises this device's shard stream once)
            e = cudaStreamSynchronize(d->stream);
            if (e == cudaSuccess && d->ws) e = cudaFree(d->ws);
            d->ws = nullptr; d->ws_bytes = 0;
            if (e == cudaSuccess) e = cudaMalloc(&d->ws, need);
            if (e == cudaSuccess) d->ws_bytes = need;
        }
        if (e == cudaSuccess) e = cudaStreamWaitEvent(d->stream, md->ready, 0);
        if (e != cudaSuccess) { rc = note_cuda_error(e); break; }
        // the shard's rows land at dst[n * stride + row_low + m]: the same entry points as a whole matrix, dst moved by row_low
        float *dst = dst_dev + p.row_low;
        rc = q6k ? q6k_mul_mat(p.w, rows, K, s->row_bytes, x_dev, N, x_row_stride_bytes, dst, dst_col_stride_bytes, d->ws, d->ws_bytes, d->stream)
                 : q4k_mul_mat(p.w, rows, K, s->row_bytes, x_dev, N, x_row_stride_bytes, dst, dst_col_stride_bytes, d->ws, d->ws_bytes, d->stream);
        if (rc != Q4K_OK) break;
        if ((e = cudaEventRecord(d->done, d->stream)) != cudaSuccess) { rc = note_cuda_error(e); break; }
        // join right away: the event's record is captured by the wait, so the next matrix may re-record it
        if ((e = cudaStreamWaitEvent((cudaStream_t)main_stream, d->done, 0)) != cudaSuccess) { rc = note_cuda_error(e); break; }
    }
    cudaSetDevice(cur);
    return rc;
}

}  // extern "C"
