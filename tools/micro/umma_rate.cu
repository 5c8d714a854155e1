// tcgen05.mma issue/throughput micro-benchmark on one SM (and on all SMs): cycles per MMA for kind::i8 / kind::f16,
// A from shared memory (SS) or tensor memory (TS), N = 64 / 128 / 256, one accumulator or alternating accumulators,
// optionally with warps draining TMEM (tcgen05.ld) at the same time.  Operands are zero bytes; only timing matters.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t saddr, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffffu) >> 4);
    d |= (uint64_t)(128 >> 4) << 16;
    d |= (uint64_t)(sbo >> 4) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

struct Cfg { int kind_f16, ts, n, alt, iters, drain; };

__global__ void __launch_bounds__(160, 1) k(Cfg c, long long *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_addr(&tmem_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_s;
    uint32_t idesc = 0;
    idesc |= (c.kind_f16 ? 1u : 2u) << 4;
    if (!c.kind_f16) idesc |= (1u << 7) | (1u << 10);
    idesc |= (uint32_t)(c.n >> 3) << 17;
    idesc |= (uint32_t)(128 >> 4) << 24;
    if (warp == 4) {
        const uint32_t tmu = __shfl_sync(0xffffffffu, tm, 0);
        const uint64_t a_desc = desc(smem_addr(smem), 1024), b_desc = desc(smem_addr(smem) + 32768, 1024);
        long long t0 = 0, t1 = 0;
        if (elect_one()) {
            t0 = clock64();
            for (int i = 0; i < c.iters; ++i) {
                const uint32_t d = tmu + (c.alt ? (uint32_t)((i & 1) * (c.n <= 128 ? 128 : 0)) : 0u);
                const uint64_t bd = b_desc + (uint64_t)((i & 3) * 16);
                if (c.ts) {
                    if (c.kind_f16) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(tmu + 384 + (uint32_t)((i & 3) * 8)), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                    else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(tmu + 384 + (uint32_t)((i & 3) * 8)), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                } else {
                    const uint64_t ad = a_desc + (uint64_t)((i & 3) * 16);
                    if (c.kind_f16) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                    else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                }
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(&bar)) : "memory");
            t1 = clock64();
        }
        __syncwarp();
        uint32_t done;
        do {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_addr(&bar)), "r"(0u) : "memory");
        } while (!done);
        const long long t2 = clock64();
        if (blockIdx.x == 0 && t0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    } else if (c.drain) {
        // four warps read accumulator columns [256, 384) over and over while the MMAs run (they never touch those)
        long long t0 = clock64();
        uint32_t s = 0;
        for (int i = 0; i < c.drain; ++i) {
            uint32_t v[8];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                         : "r"(tm + ((uint32_t)(32 * warp) << 16) + 256 + (uint32_t)((i & 15) * 8)) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            s += v[0] + v[7];
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && threadIdx.x == 0) { out[2] = t1 - t0; out[3] = s; }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

int main() {
    long long *out, h[4];
    cudaMalloc(&out, 32);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    const int iters = 512;
    for (int grid : {1, 148})
        for (int f16 = 0; f16 < 2; ++f16)
            for (int ts = 0; ts < 2; ++ts)
                for (int n : {64, 128, 256})
                    for (int alt = 0; alt < 2; ++alt)
                        for (int drain : {0, 4096}) {
                            if (alt && n == 256) continue;
                            if (drain && (alt || grid > 1)) continue;
                            Cfg c{f16, ts, n, alt, iters, drain};
                            cudaMemset(out, 0, 32);
                            k<<<grid, 160, 96 * 1024>>>(c, out);
                            cudaError_t e = cudaDeviceSynchronize();
                            cudaMemcpy(h, out, 32, cudaMemcpyDeviceToHost);
                            printf("grid %3d kind %s %s N %3d %s drain %4d : issue %6.1f cyc/mma, complete %6.1f cyc/mma = %5.0f MAC/clk/SM%s",
                                   grid, f16 ? "f16" : "i8 ", ts ? "TS" : "SS", n, alt ? "alt-D " : "same-D", drain, (double)h[0] / iters,
                                   (double)h[1] / iters, 128.0 * n * (f16 ? 16 : 32) * iters / (double)h[1], e ? " ERR" : "");
                            if (drain) printf("  | ld.x8+wait: %.1f cyc each", (double)h[2] / drain);
                            printf("\n");
                        }
    return 0;
}
