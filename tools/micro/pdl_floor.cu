// Launch-chain floor on B200: microseconds per dependent launch inside a CUDA graph, with and without
// programmatic dependent launch, for kernels that do (a) nothing, (b) a dependent global load + store,
// (c) a load, a warp reduction and a __syncthreads (the shape of the GEMV's activation staging).
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_empty(const float *x, float *y, int mode) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (mode >= 1) {
        float v = x[(threadIdx.x & 255) + (blockIdx.x & 7) * 256];
        if (mode >= 2) {
            for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
            __shared__ float s[8];
            if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
            __syncthreads();
            v = s[threadIdx.x & 7];
        }
        if (mode >= 3) {     // dependent shared-memory chain ~ the integer phase's latency
            for (int i = 0; i < 200; ++i) v = __fmaf_rn(v, 1.0001f, 0.5f);
        }
        if (threadIdx.x == 0) y[blockIdx.x] = v;
    }
}

static float run(int grid, int mode, bool pdl, int chain, int reps, int threads = 256, int smem = 0, int cluster = 1) {
    if (smem) cudaFuncSetAttribute(k_empty, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    float *x, *y;
    cudaMalloc(&x, 1 << 20); cudaMalloc(&y, 1 << 20);
    cudaMemset(x, 0, 1 << 20);
    cudaStream_t st; cudaStreamCreate(&st);
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < chain; ++i) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(threads); cfg.stream = st; cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[2];
        int na = 0;
        if (pdl) { at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[na].val.programmaticStreamSerializationAllowed = 1; ++na; }
        if (cluster > 1) { at[na].id = cudaLaunchAttributeClusterDimension; at[na].val.clusterDim.x = cluster; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1; ++na; }
        cfg.attrs = at; cfg.numAttrs = na;
        cudaLaunchKernelEx(&cfg, k_empty, (const float *)x, y, mode);
    }
    cudaStreamEndCapture(st, &g);
    cudaGraphInstantiate(&ge, g, 0);
    cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, st);
    for (int r = 0; r < reps; ++r) cudaGraphLaunch(ge, st);
    cudaEventRecord(e1, st);
    cudaStreamSynchronize(st);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaGraphExecDestroy(ge); cudaGraphDestroy(g); cudaStreamDestroy(st); cudaFree(x); cudaFree(y);
    return ms * 1e3f / (chain * reps);
}

int main() {
    for (int pdl = 0; pdl < 2; ++pdl)
        for (int grid : {1, 148, 296})
            for (int mode = 0; mode < 4; ++mode)
                printf("pdl=%d grid=%3d mode=%d  %.3f us/launch\n", pdl, grid, mode, run(grid, mode, pdl, 200, 20));
    for (int smem : {0, 40 << 10, 105 << 10, 200 << 10})
        for (int threads : {256, 320})
            for (int grid : {128, 296})
                printf("pdl=1 grid=%3d threads=%d smem=%3dKB mode=2  %.3f us/launch\n", grid, threads, smem >> 10, run(grid, 2, true, 200, 20, threads, smem));
    for (int cluster : {1, 2, 4})
        for (int grid : {128, 256})
            for (int pdl = 0; pdl < 2; ++pdl)
                printf("cluster=%d pdl=%d grid=%3d threads=320 smem=105KB mode=2  %.3f us/launch\n", cluster, pdl, grid, run(grid, 2, pdl, 200, 20, 320, 105 << 10, cluster));
    cudaError_t e = cudaDeviceSynchronize();
    printf("status %s\n", cudaGetErrorString(e));
    return 0;
}
