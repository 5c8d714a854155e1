"""Small single-shape runner for ncu: a few GEMV launches over rotating weight copies."""
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
ap.add_argument("--M", type=int, default=28672)
ap.add_argument("--K", type=int, default=8192)
ap.add_argument("--N", type=int, default=1)
ap.add_argument("--iters", type=int, default=6)
ap.add_argument("--order", type=int, default=0)
ap.add_argument("--cfg", type=int, default=0)
ap.add_argument("--q6k", action="store_true", help="Q6_K weights (gemv_q6k_kernel)")
a = ap.parse_args()
lib = q4k_b200.load()
lib.q4k_debug_set_gemv_order(a.order)
lib.q4k_debug_set_gemv_config(a.cfg)
blk = 210 if a.q6k else 144
copies = max(2, (320 << 20) // (a.M * (a.K // 256) * blk) + 1)
if a.q6k:
    from q6k_sweep import rand_q6k
    w = rand_q6k(a.M, a.K, copies)
else:
    w = rand_q4k(a.M, a.K, copies)
x = torch.randn((a.N, a.K), device="cuda")
for i in range(a.iters):
    y = q4k_b200.mul_mat_q6k(w[i % copies], x) if a.q6k else q4k_b200.mul_mat(w[i % copies], x)
torch.cuda.synchronize()
print("ok", float(y.abs().sum()))
