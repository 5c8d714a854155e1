"""Decode with 2..8 columns through q4k_mul_mat (quantizer fused or separate as the C ABI decides), CUDA-graph timed over
rotating weight copies.  Development tool."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from gemv_sweep import graph_time_us, rand_q4k  # noqa: E402

import os as _os
SHAPES = [tuple(map(int, s.split("x"))) for s in _os.environ.get("SHAPES", "5632x2048,4096x14336,14336x4096,28672x8192,8192x28672").split(",")]
for M, K in SHAPES:
    wbytes = M * (K // 256) * 144
    copies = max(2, (320 << 20) // wbytes + 1)
    w = rand_q4k(M, K, copies)
    for N in (2, 4, 8):
        x = torch.randn((N, K), device="cuda")
        dst = torch.empty((N, M), device="cuda")
        ws = torch.empty(q4k_b200.workspace_bytes(K, N), dtype=torch.uint8, device="cuda")
        us = graph_time_us(lambda i: q4k_b200.mul_mat(w[i % copies], x, dst, ws), per_graph=max(copies, 10), replays=5)
        print(dict(M=M, K=K, N=N, us=round(us, 2), gbs=round((wbytes + N * K * 4 + M * N * 4) / us / 1e3, 1)), flush=True)
    del w
