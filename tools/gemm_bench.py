"""Prefill GEMM timing on one GPU: q4k_mul_mat (planar quantizer + tcgen05 GEMM) per shape, int8 TOPS = 2MNK / time.
Development tool; bench.py carries the judged numbers."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from gemv_sweep import rand_q4k  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shapes", default="4096x14336x512,14336x4096x512,5632x2048x512,2048x5632x512,2048x2048x512,4096x14336x32")
ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()
for shp in a.shapes.split(","):
    M, K, N = map(int, shp.split("x"))
    w = rand_q4k(M, K, 1)[0]
    x = torch.randn((N, K), device="cuda")
    out = torch.empty((N, M), device="cuda")
    ws = torch.empty(q4k_b200.workspace_bytes(K, N), dtype=torch.uint8, device="cuda")
    for _ in range(3):
        q4k_b200.mul_mat(w, x, out, ws)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(True), torch.cuda.Event(True)
    s.record()
    for _ in range(a.iters):
        q4k_b200.mul_mat(w, x, out, ws)
    e.record()
    torch.cuda.synchronize()
    us = s.elapsed_time(e) * 1e3 / a.iters
    print(dict(M=M, K=K, N=N, us=round(us, 1), tops=round(2 * M * N * K / us / 1e6, 1)), flush=True)
