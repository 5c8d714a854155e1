"""The four prefill launches of a TinyLlama layer at 512 tokens (q/k/v group, attn_output, gate/up group, ffn_down) and a few
others through q4k_mul_mat_group_ws, us per launch pair (quantizer + GEMM).  Q4K_B200_LIB selects another build.  (GPU)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
import bench  # noqa: E402
import q4k_b200  # noqa: E402

dev = "cuda:0"
gen = torch.Generator(device=dev)
gen.manual_seed(1)
N = 512
for label, Ms, K in (("qkv", (2048, 256, 256), 2048), ("attn_output", (2048,), 2048), ("gate/up", (5632, 5632), 2048), ("ffn_down", (2048,), 5632),
                     ("4096x14336", (4096,), 14336), ("14336x4096", (14336,), 4096), ("70b gate/up", (28672, 28672), 8192)):
    ws = [bench.rand_q4k(torch, M, K, dev, gen) for M in Ms]
    x = torch.randn((N, K), device=dev, generator=gen)
    outs = None
    for _ in range(3):
        outs = q4k_b200.mul_mat_group_ws(ws, x)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 20
    a.record()
    for _ in range(iters):
        q4k_b200.mul_mat_group_ws(ws, x)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / iters
    ops = sum(2.0 * M * K * N for M in Ms)
    print(f"{label:14s} {Ms} K={K}: {us:8.1f} us  {ops / us / 1e6:7.1f} TOPS", flush=True)
    del ws, x, outs
    torch.cuda.empty_cache()
