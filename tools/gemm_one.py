"""One Q4_K prefill shape through q4k_mul_mat, a few calls (for ncu).  Usage: python tools/gemm_one.py M K N [iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
import bench  # noqa: E402
import q4k_b200  # noqa: E402

M, K, N = (int(a) for a in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
gen = torch.Generator(device="cuda:0")
gen.manual_seed(1)
w = bench.rand_q4k(torch, M, K, "cuda:0", gen)
x = torch.randn((N, K), device="cuda:0", generator=gen)
for _ in range(iters):
    y = q4k_b200.mul_mat(w, x)
torch.cuda.synchronize()
print("ok", float(y.abs().mean()))
