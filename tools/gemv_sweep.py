"""GEMV sweep on one GPU: achieved GB/s of algorithmic bytes per shape / batch / float order / tile config.
Launches are captured in a CUDA graph (no host launch overhead in the number) and rotate through >= 2x L2
of distinct weight copies so the figures are HBM, not L2.  Development tool."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
import q4k_b200  # noqa: E402


def rand_q4k(M, K, copies):
    nb = K // 256
    w = torch.randint(0, 256, (copies, M, nb, 144), dtype=torch.uint8, device="cuda")
    w[..., 0] = 0x00; w[..., 1] = 0x2c; w[..., 2] = 0x00; w[..., 3] = 0x28      # d = 2^-4, dmin = 2^-5 (fp16)
    return w.view(copies, M, nb * 144)


def graph_time_us(launch, per_graph, replays=10):
    """launch(i) enqueues one kernel on the current stream; returns microseconds per launch."""
    launch(0)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(per_graph):
            launch(i)
    g.replay()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(True), torch.cuda.Event(True)
    s.record()
    for _ in range(replays):
        g.replay()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) * 1e3 / (per_graph * replays)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="0")
    ap.add_argument("--ns", default="1,2,4,8")
    ap.add_argument("--orders", default="exact,fast")
    ap.add_argument("--shapes", default="5632x2048,2048x5632,4096x14336,14336x4096,28672x8192,8192x28672")
    a = ap.parse_args()
    lib = q4k_b200.load()
    out = []
    for shp in a.shapes.split(","):
        M, K = map(int, shp.split("x"))
        wbytes = M * (K // 256) * 144
        copies = max(2, (320 << 20) // wbytes + 1)
        w = rand_q4k(M, K, copies)
        for N in map(int, a.ns.split(",")):
            x = torch.randn((N, K), device="cuda")
            xq = q4k_b200.quantize_q8_K(x)
            dst = torch.empty((N, M), device="cuda")
            for order in a.orders.split(","):
                lib.q4k_debug_set_gemv_order(1 if order == "exact" else 2)
                for cfg in map(int, a.configs.split(",")):
                    lib.q4k_debug_set_gemv_config(cfg)

                    def launch(i):
                        rc = lib.q4k_mul_mat_q8(w[i % copies].data_ptr(), M, K, (K // 256) * 144, xq.data_ptr(), N, xq.stride(0),
                                                dst.data_ptr(), M * 4, torch.cuda.current_stream().cuda_stream)
                        assert rc == 0
                    us = graph_time_us(launch, per_graph=max(copies, 20))
                    algo = wbytes + N * K * 4 + M * N * 4
                    out.append(dict(M=M, K=K, N=N, order=order, cfg=cfg, us=round(us, 2), gbs=round(algo / us / 1e3, 1)))
                    print(out[-1], flush=True)
        del w
        torch.cuda.empty_cache()
    lib.q4k_debug_set_gemv_order(0)
    lib.q4k_debug_set_gemv_config(0)
    for N, K in ((1, 2048), (1, 14336), (512, 2048), (512, 14336)):
        x = torch.randn((N, K), device="cuda")
        y = torch.empty((N, K // 256 * 292), dtype=torch.uint8, device="cuda")
        us = graph_time_us(lambda i: q4k_b200.quantize_q8_K(x, y), 20)
        print(dict(quantize=(N, K), us=round(us, 2), gbs=round((N * K * 4 + y.numel()) / us / 1e3, 1)), flush=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gemv_sweep.json"), "w"))


if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    main()
