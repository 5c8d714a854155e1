"""TinyLlama-1.1B Q4_K_M prompt pass at 512 tokens as one q4k_plan: the 134 Q4_K mul_mats (the bench's tinyllama_pp512) and all 155
(Q4_K + Q6_K).  Usage: python tools/pp512_bench.py [--iters 5]   (GPU)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
import bench  # noqa: E402
import q4k_b200  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=5)
args = ap.parse_args()
dev = "cuda:0"
gen = torch.Generator(device=dev)
gen.manual_seed(9)
N = 512
for label, mats in (("134 Q4_K", [(n, M, K, False) for n, M, K in bench.tinyllama_token_matmuls()]), ("155 Q4_K + Q6_K", bench.tinyllama_full_token_matmuls())):
    ids = bench.shared_input_ids([(n, M, K) for n, M, K, _ in mats])
    plan = q4k_b200.GraphPlan(0)
    keep = []
    for (name, M, K, q6), iid in zip(mats, ids):
        w = (bench.rand_q6k if q6 else bench.rand_q4k)(torch, M, K, dev, gen)
        keep.append(w)
        plan.add(w, N, iid, q6k=q6)
    plan.finalize()
    for iid in set(ids):
        plan.input(iid).normal_(generator=gen)
    t = bench.timed_steps(torch, plan.run, args.iters, 2, lambda: None) / args.iters
    ops = sum(2.0 * M * K * N for _, M, K, _ in mats)
    print(f"pp512 {label}: {t:.3f} ms per pass, {ops / t / 1e9:.1f} TOPS, {512 / t * 1e3:.0f} tok/s, {int(plan.launches)} launches", flush=True)
    del plan, keep
    torch.cuda.empty_cache()
