// How fast can bulk copies (UBLKCP) alone pull a matrix out of HBM with the decode kernel's access pattern?
// Producer warp + slot ring exactly as gemv_q4k_mma.cu, consumers only wait and release.  Variants: piece size / slots / CTAs per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// rows x row_bytes matrix; a tile = 16 rows; a round = `piece` bytes of each of the 16 rows (piece = 1152) or, contig = 1,
// one 16 * piece contiguous copy (tile bytes consumed front to back)
__global__ void __launch_bounds__(64) k(const uint8_t *w, int64_t tiles, int row_bytes, int piece, int S, int contig, unsigned *sink) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem), *empty = full + 8;
    uint8_t *slots = smem + 128;
    const int slot_bytes = 16 * (piece + 16);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full[s]), 1); mbar_init(smem_u32(&empty[s]), 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int rounds = row_bytes / piece;
    if (warp == 1) {
        int s = 0; uint32_t ep = 1;
        for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            const uint8_t *base = w + tile * 16 * (int64_t)row_bytes;
            for (int rd = 0; rd < rounds; ++rd) {
                mbar_wait(smem_u32(&empty[s]), ep);
                const uint32_t bar = smem_u32(&full[s]);
                if (lane == 0) mbar_expect_tx(bar, 16 * piece);
                __syncwarp();
                if (contig) { if (lane == 0) bulk_g2s(smem_u32(slots + s * slot_bytes), base + (int64_t)rd * 16 * piece, 16 * piece, bar); }
                else if (lane < 16) bulk_g2s(smem_u32(slots + s * slot_bytes + lane * (piece + 16)), base + lane * (int64_t)row_bytes + (int64_t)rd * piece, piece, bar);
                if (++s == S) { s = 0; ep ^= 1u; }
            }
        }
        return;
    }
    int s = 0; uint32_t par = 0; unsigned acc = 0;
    for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x)
        for (int rd = 0; rd < rounds; ++rd) {
            mbar_wait(smem_u32(&full[s]), par);
            acc += *reinterpret_cast<const unsigned *>(slots + s * slot_bytes + lane * 16);
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&empty[s]));
            if (++s == S) { s = 0; par ^= 1u; }
        }
    if (acc == 0x12345678u) *sink = acc;
}

int main() {
    const int64_t rows = 28672; const int row_bytes = 4608;             // 28672 x 8192 Q4_K
    const size_t bytes = (size_t)rows * row_bytes;
    const int copies = 3;
    uint8_t *w; unsigned *sink;
    cudaMalloc(&w, bytes * copies); cudaMalloc(&sink, 4);
    cudaMemset(w, 1, bytes * copies);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 << 10);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int contig = 0; contig < 2; ++contig)
        for (int piece : {1152, 2304, 4608})
            for (int per_sm : {1, 2, 4})
                for (int S : {2, 4, 6}) {
                    const int smem = 128 + S * 16 * (piece + 16);
                    if ((size_t)smem * per_sm > (220u << 10)) continue;
                    const int grid = 148 * per_sm;
                    for (int i = 0; i < 2; ++i) k<<<grid, 64, smem>>>(w + (i % copies) * bytes, rows / 16, row_bytes, piece, S, contig, sink);
                    cudaEventRecord(e0);
                    const int reps = 9;
                    for (int i = 0; i < reps; ++i) k<<<grid, 64, smem>>>(w + (i % copies) * bytes, rows / 16, row_bytes, piece, S, contig, sink);
                    cudaEventRecord(e1); cudaEventSynchronize(e1);
                    float ms; cudaEventElapsedTime(&ms, e0, e1);
                    printf("contig=%d piece=%4d per_sm=%d S=%d in-flight/SM=%3d KB: %7.1f GB/s  (%s)\n", contig, piece, per_sm, S,
                           per_sm * S * 16 * piece >> 10, bytes * reps / (ms * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
                }
    return 0;
}
