// Per-phase clock stamps of the MMA decode kernel inside a PDL launch chain (development).  Single TU: includes the
// kernel source with -DQ4K_TIMING and stubs the two symbols it takes from c_api.cu.
#include <cstdio>
#include <vector>
#include "../../ggml-neon-opt_b200/csrc/gemv_q4k_mma.cu"

namespace q4k {
void count_launch() {}
int sm_count() { return 148; }
}

int main(int argc, char **argv) {
    const int M = argc > 1 ? atoi(argv[1]) : 2048, K = argc > 2 ? atoi(argv[2]) : 2048, N = 1, chain = 40;
    const size_t wbytes = (size_t)M * (K / 256) * 144;
    const int copies = (int)((400u << 20) / wbytes) + 1 > 64 ? 64 : (int)((400u << 20) / wbytes) + 1;
    uint8_t *w; float *x, *dst;
    cudaMalloc(&w, wbytes * copies); cudaMalloc(&x, K * 4); cudaMalloc(&dst, M * 4);
    std::vector<uint8_t> hw(wbytes);
    for (size_t i = 0; i < wbytes; ++i) hw[i] = (uint8_t)(i * 2654435761u >> 13);
    for (size_t b = 0; b < wbytes; b += 144) { hw[b] = 0; hw[b + 1] = 0x2c; hw[b + 2] = 0; hw[b + 3] = 0x28; }
    for (int c = 0; c < copies; ++c) cudaMemcpy(w + c * wbytes, hw.data(), wbytes, cudaMemcpyHostToDevice);
    std::vector<float> hx(K);
    for (int i = 0; i < K; ++i) hx[i] = (float)((i * 37) % 101 - 50) / 17.0f;
    cudaMemcpy(x, hx.data(), K * 4, cudaMemcpyHostToDevice);
    cudaStream_t st; cudaStreamCreate(&st);
    auto launch = [&](int i) {
        const void *wv[1] = {w + (size_t)(i % copies) * wbytes};
        const int64_t Mv[1] = {M};
        const float *xv[1] = {x};
        float *dv[1] = {dst};
        return q4k::launch_gemv_q4k_mma(1, wv, Mv, K, (K / 256) * 144, nullptr, xv, N, K * 4, dv, M * 4, nullptr, 0, 0, true, st);
    };
    launch(0); cudaStreamSynchronize(st);
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < chain; ++i) launch(i);
    cudaStreamEndCapture(st, &g);
    cudaGraphInstantiate(&ge, g, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
    cudaEventRecord(e0, st);
    for (int r = 0; r < 10; ++r) cudaGraphLaunch(ge, st);
    cudaEventRecord(e1, st); cudaStreamSynchronize(st);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long t[32];
    cudaMemcpyFromSymbol(t, ::g_ticks, sizeof(t));
    printf("M=%d K=%d: %.3f us/launch (%s)\n", M, K, ms * 1e3 / (chain * 10), cudaGetErrorString(cudaGetLastError()));
    const char *names[] = {"entry", "init+sync", "griddep_wait", "stage", "stage barrier", "first full wait", "compute tile 0", "last tile done", "epilogue barrier", "producer first issue"};
    for (int i = 1; i <= 7; ++i) printf("  %-20s +%6lld cycles (at %6lld)\n", names[i], t[i] - t[i - 1], t[i] - t[0]);
    printf("  %-20s at %6lld cycles\n", names[9], t[9] - t[0]);
    printf("  barrier -> before full wait %lld cycles, the wait itself %lld\n", t[14] - t[4], t[5] - t[14]);
    printf("  stage: loads landed at +%lld cycles after wait-return, quantize +%lld\n", t[13] - t[2], t[3] - t[13]);
    printf("  quantize (warp 0, last block): lane scan +%lld, butterfly +%lld, ballots +%lld, fdiv/round/pack +%lld, sums+d +%lld, stores +%lld\n",
           t[16] - t[13], t[17] - t[16], t[18] - t[17], t[19] - t[18], t[20] - t[19], t[3] - t[20]);
    printf("  reducer (first tile): own segments done at %lld, remote parts landed +%lld, stored +%lld (cycles after wait-return)\n",
           t[21] - t[2], t[22] - t[21], t[23] - (t[22] > t[21] ? t[22] : t[21]));
    printf("  globaltimer: entry->wait-return %lld ns, wait-return->end %lld ns\n", t[11] - t[10], t[12] - t[11]);
    return 0;
}
