"""Q6_K decode GEMV sweep on one GPU: achieved GB/s of algorithmic bytes (M*K/256*210 + N*K*4 + M*N*4) per shape and
batch.  Launches are captured in a CUDA graph and rotate through >= 2.5x L2 of distinct weight copies.  Development tool;
Q6K_DBG_STAGES=<n> overrides the ring depth."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from gemv_sweep import graph_time_us  # noqa: E402


def rand_q6k(M, K, copies):
    nb = K // 256
    w = torch.randint(0, 256, (copies, M, nb, 210), dtype=torch.uint8, device="cuda")
    w[..., 208] = 0x00
    w[..., 209] = 0x1c        # d = 2^-8 (fp16)
    return w.view(copies, M, nb * 210)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ns", default="1,2,4,8")
    ap.add_argument("--shapes", default="256x2048,2048x5632,32000x2048,4096x14336,14336x4096,28672x8192,8192x28672")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    out = []
    for shp in a.shapes.split(","):
        M, K = map(int, shp.split("x"))
        wbytes = M * (K // 256) * 210
        copies = max(2, (320 << 20) // wbytes + 1)
        w = rand_q6k(M, K, copies)
        for N in map(int, a.ns.split(",")):
            x = torch.randn((N, K), device="cuda")
            dst = torch.empty((N, M), device="cuda")
            ws = torch.empty(max(16, N * (K // 256) * 292), dtype=torch.uint8, device="cuda")
            us = graph_time_us(lambda i: q4k_b200.mul_mat_q6k(w[i % copies], x, dst, ws), per_graph=max(copies, 20))
            gbs = (wbytes + N * K * 4 + M * N * 4) / us / 1e3
            row = dict(M=M, K=K, N=N, us=round(us, 2), gbs=round(gbs, 1), copies=copies)
            print(row, flush=True)
            out.append(row)
        del w
    if a.out:
        json.dump(out, open(a.out, "w"), indent=0)


if __name__ == "__main__":
    main()
