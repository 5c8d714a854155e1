// plan.cu — graph plans for the Q4_K mul_mat path (q4k_plan_* of include/q4k_b200.h).
//
// The reference's backend interface has optional graph plans (graph_plan_create / graph_plan_compute,
// L/ggml/src/ggml-backend-impl.h:100-108): "compute graph with a plan".  This is that idea for a sequence of Q4_K
// mul_mats: record once, replay as one CUDA graph.  What the plan adds over calling q4k_mul_mat in a loop:
//   * the launches are captured with their programmatic-dependent-launch edges, so a replay costs one graph launch
//     and the weight prefetch of kernel i + 1 overlaps the tail of kernel i;
//   * consecutive mul_mats that read the same activations (q/k/v, gate/up) are fused into one launch;
//   * inputs and outputs are packed into two device buffers, so a host caller does ONE H2D and ONE D2H per replay
//     instead of one pair per mul_mat.
#include "../../include/q4k_b200.h"

#include <cstdio>
#include <vector>

#include "q4k_kernels.h"

namespace q4k {
int note_cuda_error(cudaError_t e);      // c_api.cu
bool gemv_exact_policy(int N);           // c_api.cu
}

struct q4k_plan {
    struct Item {
        const void *w;
        int64_t M, K, w_stride, N;
        int input_id;
        size_t out_off;
        int q6;               // 1: block_q6_K weights (q6k_mul_mat), 0: block_q4_K
    };
    struct Input {
        int id;
        int64_t N, K;
        size_t off;
    };
    int device = 0;
    std::vector<Item> items;
    std::vector<Input> inputs;
    size_t in_bytes = 0, out_bytes = 0, ws_bytes = 0;
    char *in_dev = nullptr, *out_dev = nullptr, *ws_dev = nullptr;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int64_t launches = 0;
    bool finalized = false;
    // host-buffer replay (q4k_plan_run_host) as a copy / compute pipeline: the items cut into parts, one CUDA graph per part;
    // part p's inputs go up and part p - 1's results come down on a side stream while part p computes
    struct Part {
        size_t begin, end;                 // items [begin, end)
        size_t in_off, in_len, out_off, out_len;
        cudaGraph_t graph;
        cudaGraphExec_t exec;
    };
    std::vector<Part> parts;
    cudaStream_t side = nullptr;
    cudaEvent_t ev_start = nullptr, ev_join = nullptr;
    std::vector<cudaEvent_t> ev_in, ev_out;

    const Input *find_input(int id) const {
        for (const auto &in : inputs) if (in.id == id) return &in;
        return nullptr;
    }
};

namespace {

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// enqueue the launches of items [begin, end) on `st` (used under stream capture)
int enqueue_range(q4k_plan *p, cudaStream_t st, size_t begin, size_t end) {
    size_t i = begin;
    while (i < end) {
        const q4k_plan::Item &a = p->items[i];
        size_t j = i + 1;
        if (!a.q6) {          // decode: the group launch of the decode kernel; prefill: one GEMM launch over all row tiles
            while (j < end && j - i < 4 && !p->items[j].q6 && p->items[j].input_id == a.input_id && p->items[j].K == a.K &&
                   p->items[j].N == a.N && p->items[j].w_stride == a.w_stride) ++j;
        }
        const q4k_plan::Input *in = p->find_input(a.input_id);
        const float *x = reinterpret_cast<const float *>(p->in_dev + in->off);
        int rc = Q4K_ERR_UNSUPPORTED;
        if (j - i > 1) {
            const void *w[4];
            int64_t M[4];
            const float *xs[4];
            float *dst[4];
            for (size_t k = i; k < j; ++k) {
                w[k - i] = p->items[k].w; M[k - i] = p->items[k].M; xs[k - i] = x;
                dst[k - i] = reinterpret_cast<float *>(p->out_dev + p->items[k].out_off);
            }
            rc = q4k_mul_mat_group_ws((int)(j - i), w, M, a.K, a.w_stride, x, a.N, a.K * 4, dst, p->ws_dev, p->ws_bytes, st);
            if (rc != Q4K_OK && rc != Q4K_ERR_UNSUPPORTED) return rc;
        }
        if (rc == Q4K_ERR_UNSUPPORTED) {            // not fusable (or a single mul_mat): one q4k_mul_mat each
            for (size_t k = i; k < j; ++k) {
                const q4k_plan::Item &b = p->items[k];
                rc = b.q6 ? q6k_mul_mat(b.w, b.M, b.K, b.w_stride, x, b.N, b.K * 4, reinterpret_cast<float *>(p->out_dev + b.out_off), b.M * 4,
                                        p->ws_dev, p->ws_bytes, st)
                          : q4k_mul_mat(b.w, b.M, b.K, b.w_stride, x, b.N, b.K * 4, reinterpret_cast<float *>(p->out_dev + b.out_off), b.M * 4,
                                        p->ws_dev, p->ws_bytes, st);
                if (rc != Q4K_OK) return rc;
            }
        }
        i = j;
    }
    return Q4K_OK;
}

int enqueue_all(q4k_plan *p, cudaStream_t st) { return enqueue_range(p, st, 0, p->items.size()); }

// Cut the items into up to `want` parts of about equal result bytes, only where the input changes (a fused group is never
// split), capture one graph per part and note the byte ranges of the packed buffers each part owns: results are packed in
// item order, inputs in order of first use, so both are contiguous per part.
int build_parts(q4k_plan *p, cudaStream_t cs, int want) {
    const size_t n = p->items.size();
    if (n < 8 || want < 2) return Q4K_OK;
    // four parts of about equal result bytes (a small first / last part — less exposed copy — measured no better: the host
    // then cannot enqueue the second graph before the first has finished)
    static const double frac[4] = {0.25, 0.5, 0.75, 2.0};
    (void)want;
    std::vector<size_t> cuts{0};
    size_t acc = 0;
    for (size_t i = 0; i < n; ++i) {
        if (i > 0 && p->items[i].input_id != p->items[i - 1].input_id && cuts.size() <= 4 &&
            (double)acc >= frac[cuts.size() - 1] * (double)p->out_bytes) cuts.push_back(i);
        acc += ((size_t)p->items[i].N * p->items[i].M * 4 + 255) & ~(size_t)255;
    }
    cuts.push_back(n);
    if (cuts.size() < 3) return Q4K_OK;
    std::vector<char> seen(p->inputs.size(), 0);
    for (size_t c = 0; c + 1 < cuts.size(); ++c) {
        q4k_plan::Part part{};
        part.begin = cuts[c]; part.end = cuts[c + 1];
        part.out_off = p->items[part.begin].out_off;
        part.out_len = (part.end < n ? p->items[part.end].out_off : p->out_bytes) - part.out_off;
        size_t lo = (size_t)-1, hi = 0;
        for (size_t i = part.begin; i < part.end; ++i) {
            for (size_t k = 0; k < p->inputs.size(); ++k) {
                if (p->inputs[k].id != p->items[i].input_id || seen[k]) continue;
                seen[k] = 1;
                const size_t o = p->inputs[k].off, e = o + (((size_t)p->inputs[k].N * p->inputs[k].K * 4 + 255) & ~(size_t)255);
                if (o < lo) lo = o;
                if (e > hi) hi = e;
            }
        }
        part.in_off = lo == (size_t)-1 ? 0 : lo;
        part.in_len = lo == (size_t)-1 ? 0 : hi - lo;
        cudaError_t e = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
        if (e != cudaSuccess) return q4k::note_cuda_error(e);
        const int rc = enqueue_range(p, cs, part.begin, part.end);
        e = cudaStreamEndCapture(cs, &part.graph);
        if (rc != Q4K_OK || e != cudaSuccess) return rc != Q4K_OK ? rc : q4k::note_cuda_error(e);
        if ((e = cudaGraphInstantiate(&part.exec, part.graph, 0)) != cudaSuccess) return q4k::note_cuda_error(e);
        p->parts.push_back(part);
    }
    cudaError_t e = cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking);
    if (e != cudaSuccess) return q4k::note_cuda_error(e);
    cudaEventCreateWithFlags(&p->ev_start, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming);
    p->ev_in.resize(p->parts.size());
    p->ev_out.resize(p->parts.size());
    for (size_t c = 0; c < p->parts.size(); ++c) {
        cudaEventCreateWithFlags(&p->ev_in[c], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&p->ev_out[c], cudaEventDisableTiming);
    }
    return Q4K_OK;
}

}  // namespace

extern "C" {

q4k_plan *q4k_plan_create(int device) {
    if (!q4k_device_ok(device)) return nullptr;
    q4k_plan *p = new q4k_plan;
    p->device = device;
    return p;
}

static int plan_add(q4k_plan *p, const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes, int64_t N, int input_id, int q6) {
    if (!p || p->finalized) return Q4K_ERR_UNSUPPORTED;
    if (!w_dev) return Q4K_ERR_BAD_POINTER;
    if (M <= 0 || N <= 0 || K <= 0 || K % Q4K_QK_K != 0) return Q4K_ERR_BAD_SHAPE;
    if (q6 ? (w_row_stride_bytes < K / Q4K_QK_K * Q4K_BLOCK_Q6K || (w_row_stride_bytes & 1))
           : (w_row_stride_bytes < K / Q4K_QK_K * Q4K_BLOCK_Q4K || (w_row_stride_bytes & 15)))
        return Q4K_ERR_BAD_SHAPE;
    const q4k_plan::Input *in = p->find_input(input_id);
    if (in && (in->N != N || in->K != K)) return Q4K_ERR_BAD_SHAPE;        // one input_id = one [N][K] buffer
    if (!in) {
        p->inputs.push_back({input_id, N, K, p->in_bytes});
        p->in_bytes += align256((size_t)N * K * 4);
    }
    p->items.push_back({w_dev, M, K, w_row_stride_bytes, N, input_id, p->out_bytes, q6});
    p->out_bytes += align256((size_t)N * M * 4);
    const size_t ws = q6 ? q6k_mul_mat_workspace_bytes(K, N) : q4k_mul_mat_workspace_bytes(K, N);
    if (ws > p->ws_bytes) p->ws_bytes = ws;
    return (int)p->items.size() - 1;
}

int q4k_plan_add_mul_mat(q4k_plan *p, const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes, int64_t N, int input_id) {
    return plan_add(p, w_dev, M, K, w_row_stride_bytes, N, input_id, 0);
}

int q4k_plan_add_mul_mat_q6k(q4k_plan *p, const void *w_dev, int64_t M, int64_t K, int64_t w_row_stride_bytes, int64_t N, int input_id) {
    return plan_add(p, w_dev, M, K, w_row_stride_bytes, N, input_id, 1);
}

int q4k_plan_finalize(q4k_plan *p) {
    if (!p || p->finalized || p->items.empty()) return Q4K_ERR_UNSUPPORTED;
    cudaError_t e = cudaSetDevice(p->device);
    if (e != cudaSuccess) return q4k::note_cuda_error(e);
    if ((e = cudaMalloc(&p->in_dev, p->in_bytes)) != cudaSuccess) return q4k::note_cuda_error(e);
    if ((e = cudaMalloc(&p->out_dev, p->out_bytes)) != cudaSuccess) return q4k::note_cuda_error(e);
    if ((e = cudaMalloc(&p->ws_dev, p->ws_bytes ? p->ws_bytes : 16)) != cudaSuccess) return q4k::note_cuda_error(e);
    if ((e = cudaMemset(p->in_dev, 0, p->in_bytes)) != cudaSuccess) return q4k::note_cuda_error(e);
    cudaStream_t cs = nullptr;
    if ((e = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking)) != cudaSuccess) return q4k::note_cuda_error(e);
    // one eager pass first: per-kernel attributes (dynamic shared memory opt-in) are set outside the capture
    int rc = enqueue_all(p, cs);
    if (rc == Q4K_OK && (e = cudaStreamSynchronize(cs)) != cudaSuccess) rc = q4k::note_cuda_error(e);
    if (rc != Q4K_OK) { cudaStreamDestroy(cs); return rc; }
    const int64_t before = q4k::launch_count();
    if ((e = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal)) != cudaSuccess) { cudaStreamDestroy(cs); return q4k::note_cuda_error(e); }
    rc = enqueue_all(p, cs);
    e = cudaStreamEndCapture(cs, &p->graph);
    p->launches = q4k::launch_count() - before;
    if (rc != Q4K_OK || e != cudaSuccess) { cudaStreamDestroy(cs); return rc != Q4K_OK ? rc : q4k::note_cuda_error(e); }
    e = cudaGraphInstantiate(&p->exec, p->graph, 0);
    if (e != cudaSuccess) { cudaStreamDestroy(cs); return q4k::note_cuda_error(e); }
    const int before_parts = (int)q4k::launch_count();
    rc = build_parts(p, cs, 4);
    (void)before_parts;
    cudaStreamDestroy(cs);
    if (rc != Q4K_OK) return rc;
    p->finalized = true;
    return Q4K_OK;
}

size_t q4k_plan_input_bytes(const q4k_plan *p) { return p ? p->in_bytes : 0; }
size_t q4k_plan_output_bytes(const q4k_plan *p) { return p ? p->out_bytes : 0; }
int64_t q4k_plan_input_offset(const q4k_plan *p, int input_id) {
    const q4k_plan::Input *in = p ? p->find_input(input_id) : nullptr;
    return in ? (int64_t)in->off : -1;
}
int64_t q4k_plan_output_offset(const q4k_plan *p, int index) {
    return (p && index >= 0 && (size_t)index < p->items.size()) ? (int64_t)p->items[index].out_off : -1;
}
void *q4k_plan_input_dev(q4k_plan *p) { return p ? p->in_dev : nullptr; }
void *q4k_plan_output_dev(q4k_plan *p) { return p ? p->out_dev : nullptr; }
int64_t q4k_plan_launches(const q4k_plan *p) { return p ? p->launches : 0; }

int q4k_plan_run(q4k_plan *p, void *stream) {
    if (!p || !p->finalized) return Q4K_ERR_UNSUPPORTED;
    const cudaError_t e = cudaGraphLaunch(p->exec, (cudaStream_t)stream);
    if (e != cudaSuccess) return q4k::note_cuda_error(e);
    q4k::add_launches(p->launches);
    return Q4K_OK;
}

int q4k_plan_run_host(q4k_plan *p, const void *in_host, void *out_host, void *stream) {
    if (!p || !p->finalized) return Q4K_ERR_UNSUPPORTED;
    if (!in_host || !out_host) return Q4K_ERR_BAD_POINTER;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e;
    if (p->parts.empty()) {                  // short plans: one copy in, the graph, one copy out
        e = cudaMemcpyAsync(p->in_dev, in_host, p->in_bytes, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return q4k::note_cuda_error(e);
        const int rc = q4k_plan_run(p, stream);
        if (rc != Q4K_OK) return rc;
        e = cudaMemcpyAsync(out_host, p->out_dev, p->out_bytes, cudaMemcpyDeviceToHost, st);
        return e == cudaSuccess ? Q4K_OK : q4k::note_cuda_error(e);
    }
    // Copy / compute pipeline.  Caller's stream: inputs of part 0, graph 0, graph 1 ...; side stream: the inputs of parts
    // 1.. right away, then the results of each part as soon as its graph has finished — the copy engines run under the
    // kernels, only the first upload and the last download stay exposed.  The caller's stream joins the side stream at the end.
    const char *in = static_cast<const char *>(in_host);
    char *out = static_cast<char *>(out_host);
    const size_t np = p->parts.size();
#define Q4K_TRY(call) do { e = (call); if (e != cudaSuccess) return q4k::note_cuda_error(e); } while (0)
    Q4K_TRY(cudaEventRecord(p->ev_start, st));                               // orders the side stream after the caller's earlier work
    if (p->parts[0].in_len) Q4K_TRY(cudaMemcpyAsync(p->in_dev + p->parts[0].in_off, in + p->parts[0].in_off, p->parts[0].in_len, cudaMemcpyHostToDevice, st));
    Q4K_TRY(cudaGraphLaunch(p->parts[0].exec, st));
    Q4K_TRY(cudaEventRecord(p->ev_out[0], st));
    Q4K_TRY(cudaStreamWaitEvent(p->side, p->ev_start, 0));
    for (size_t c = 1; c < np; ++c) {
        const q4k_plan::Part &pt = p->parts[c];
        if (pt.in_len) Q4K_TRY(cudaMemcpyAsync(p->in_dev + pt.in_off, in + pt.in_off, pt.in_len, cudaMemcpyHostToDevice, p->side));
        Q4K_TRY(cudaEventRecord(p->ev_in[c], p->side));
    }
    for (size_t c = 1; c < np; ++c) {
        Q4K_TRY(cudaStreamWaitEvent(st, p->ev_in[c], 0));
        Q4K_TRY(cudaGraphLaunch(p->parts[c].exec, st));
        Q4K_TRY(cudaEventRecord(p->ev_out[c], st));
    }
    for (size_t c = 0; c < np; ++c) {
        const q4k_plan::Part &pt = p->parts[c];
        Q4K_TRY(cudaStreamWaitEvent(p->side, p->ev_out[c], 0));
        Q4K_TRY(cudaMemcpyAsync(out + pt.out_off, p->out_dev + pt.out_off, pt.out_len, cudaMemcpyDeviceToHost, p->side));
    }
    Q4K_TRY(cudaEventRecord(p->ev_join, p->side));
    Q4K_TRY(cudaStreamWaitEvent(st, p->ev_join, 0));
#undef Q4K_TRY
    q4k::add_launches(p->launches);
    return Q4K_OK;
}

void q4k_plan_destroy(q4k_plan *p) {
    if (!p) return;
    cudaSetDevice(p->device);
    if (p->exec) cudaGraphExecDestroy(p->exec);
    if (p->graph) cudaGraphDestroy(p->graph);
    for (auto &pt : p->parts) { if (pt.exec) cudaGraphExecDestroy(pt.exec); if (pt.graph) cudaGraphDestroy(pt.graph); }
    for (auto ev : p->ev_in) if (ev) cudaEventDestroy(ev);
    for (auto ev : p->ev_out) if (ev) cudaEventDestroy(ev);
    if (p->ev_start) cudaEventDestroy(p->ev_start);
    if (p->ev_join) cudaEventDestroy(p->ev_join);
    if (p->side) cudaStreamDestroy(p->side);
    cudaFree(p->in_dev);
    cudaFree(p->out_dev);
    cudaFree(p->ws_dev);
    delete p;
}

}  // extern "C"
