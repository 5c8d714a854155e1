// peer_sync.cu — the cross-GPU step of the row-sharded decode path (q4k_peer_signal_wait of include/q4k_b200.h).
//
// q4k_mul_mat_peers makes the kernel's epilogue the all-gather: every rank stores its result rows straight into the
// output buffer of every peer over NVLink.  What is left of a collective is "have all ranks' stores landed?" — one
// flag per (rank, peer) in peer-mapped memory.  This kernel (one CTA) bumps a device-side epoch (so that a CUDA graph
// can replay it), publishes it with a system-scope release store into every peer's flag row, and spins with
// system-scope acquire loads until every peer's epoch has arrived in its own.  The stores of the preceding kernel are
// ordered before the flags by the kernel boundary + the cumulative system fence.  No NCCL, no host round trip.
//
// The reference has no such step: ggml-cuda's row split gathers dst slices to the main device with
// cudaMemcpyPeerAsync + events (L/ggml/src/ggml-cuda/ggml-cuda.cu:1720-1760).
#include "../../include/q4k_b200.h"

#include <cuda_runtime.h>
#include <stdint.h>

#include "q4k_kernels.h"

namespace q4k {
int note_cuda_error(cudaError_t e);      // c_api.cu

namespace {

struct PeerSyncParams {
    uint32_t *local_flags;        // [world] in this rank's peer-mapped buffer: slot r is written by rank r
    uint32_t *peer_flags[8];      // the same buffer on every rank (peer-mapped pointers), indexed by rank
    uint32_t *epoch;              // this rank's call counter (plain device memory)
    int world, rank;
};

__global__ void __launch_bounds__(32) peer_signal_wait_kernel(const PeerSyncParams p) {
    // programmatic dependent launch on both sides: this CTA is already resident when the mul_mat before it finishes, and
    // the mul_mat after it starts prefetching its weights while the flags are in flight (its own griddepcontrol.wait
    // still holds its activation reads back until this kernel has completed, i.e. until every peer's stores have landed)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int t = threadIdx.x;
    uint32_t e = 0;
    if (t == 0) {
        e = *p.epoch + 1u;
        *p.epoch = e;
    }
    e = __shfl_sync(0xffffffffu, e, 0);
    if (t < p.world && t != p.rank) {
        // release.sys is cumulative: the previous kernel's peer stores (ordered before this thread by the kernel boundary)
        // are visible at the peer before the flag is
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.peer_flags[t] + p.rank), "r"(e) : "memory");
        uint32_t seen;
        do {
            asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.local_flags + t) : "memory");
        } while ((int32_t)(seen - e) < 0);          // wrap-safe "seen >= e"
    }
    __syncwarp();
    asm volatile("fence.acq_rel.sys;" ::: "memory");   // acquire side: results read after this kernel see the peers' stores
}

}  // namespace
}  // namespace q4k

extern "C" int q4k_peer_signal_wait(void *const *flags_peers, int world, int rank, void *epoch_dev, void *stream) {
    if (world < 1 || world > 8 || rank < 0 || rank >= world) return Q4K_ERR_BAD_SHAPE;
    if (!flags_peers || !epoch_dev) return Q4K_ERR_BAD_POINTER;
    q4k::PeerSyncParams p;
    for (int r = 0; r < 8; ++r) p.peer_flags[r] = nullptr;
    for (int r = 0; r < world; ++r) {
        if (!flags_peers[r]) return Q4K_ERR_BAD_POINTER;
        p.peer_flags[r] = static_cast<uint32_t *>(flags_peers[r]);
    }
    p.local_flags = p.peer_flags[rank];
    p.epoch = static_cast<uint32_t *>(epoch_dev);
    p.world = world; p.rank = rank;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1);
    cfg.blockDim = dim3(32);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = q4k::knobs().pdl ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, q4k::peer_signal_wait_kernel, p);
    q4k::count_launch();
    return e == cudaSuccess ? Q4K_OK : q4k::note_cuda_error(e);
}
