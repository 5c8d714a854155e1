// Latency / throughput of one quantize_block_q8_K per warp: REDUX version (q4k_common.cuh) vs the shuffle butterfly.
#include <cstdio>
#include "../../ggml-neon-opt_b200/csrc/q4k_common.cuh"
using namespace q4k;

__device__ __forceinline__ float quant_shfl(const float (&v)[8], int lane, uint32_t &w_lo, uint32_t &w_hi, int &s_lo, int &s_hi) {
    float amax = 0.0f, smax = 0.0f;
    int idx = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float ax = fabsf(v[i]);
        if (ax > amax) { amax = ax; smax = v[i]; idx = (i < 4 ? 4 * lane + i : 128 + 4 * lane + (i - 4)); }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, amax, off);
        const float os = __shfl_xor_sync(0xffffffffu, smax, off);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, off);
        if (oa > amax || (oa == amax && oi < idx)) { amax = oa; smax = os; idx = oi; }
    }
    int q[8];
    float d_out = 0.0f;
    if (amax == 0.0f) {
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = 0;
    } else {
        const float iscale = __fdiv_rn(-127.0f, smax);
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = min(127, nearest_int(__fmul_rn(iscale, v[i])));
        d_out = __fdiv_rn(1.0f, iscale);
    }
    w_lo = (uint32_t)(q[0] & 0xff) | ((uint32_t)(q[1] & 0xff) << 8) | ((uint32_t)(q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
    w_hi = (uint32_t)(q[4] & 0xff) | ((uint32_t)(q[5] & 0xff) << 8) | ((uint32_t)(q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
    s_lo = q[0] + q[1] + q[2] + q[3];
    s_hi = q[4] + q[5] + q[6] + q[7];
    s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 1); s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 1);
    s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 2); s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 2);
    s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 4); s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 4);
    return d_out;
}

template <int MODE>
__global__ void k(const float *x, float *out, long long *cyc, int iters) {
    const int lane = threadIdx.x & 31;
    const float4 a = reinterpret_cast<const float4 *>(x)[lane], c = reinterpret_cast<const float4 *>(x)[32 + lane];
    float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
    float acc = 0.0f;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        uint32_t w_lo, w_hi; int s_lo, s_hi;
        const float d = MODE == 0 ? quantize_block_q8_K<8>(v, lane, w_lo, w_hi, s_lo, s_hi) : quant_shfl(v, lane, w_lo, w_hi, s_lo, s_hi);
        acc += d + (float)(w_lo ^ w_hi) * 1e-30f + (float)(s_lo + s_hi) * 1e-30f;
        v[i & 7] += acc * 1e-20f;          // serialise the iterations
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
    float *x, *out; long long *cyc, h;
    cudaMalloc(&x, 4096); cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 8);
    float hx[256];
    for (int i = 0; i < 256; ++i) hx[i] = (float)((i * 37) % 101 - 50) / 17.0f;
    cudaMemcpy(x, hx, sizeof(hx), cudaMemcpyHostToDevice);
    const int iters = 64;
    for (int warps : {1, 4, 8, 16}) {
        k<0><<<148, warps * 32>>>(x, out, cyc, iters); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("redux  warps/SM=%2d: %6.1f cycles per block\n", warps, (double)h / iters);
        k<1><<<148, warps * 32>>>(x, out, cyc, iters); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("shfl   warps/SM=%2d: %6.1f cycles per block\n", warps, (double)h / iters);
    }
    printf("status %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
