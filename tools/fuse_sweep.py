"""Decode N = 1: quantizer fused into the GEMV launch (every CTA quantizes the activations itself) against a separate
quantize kernel + GEMV on Q8_K, per K.  Where is the crossover for the fuse limits?  Development tool."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from gemv_sweep import graph_time_us, rand_q4k  # noqa: E402
from q6k_sweep import rand_q6k  # noqa: E402

lib = q4k_b200.load()
for q6 in (False, True):
    for M, K in [(8192, 2048), (8192, 4096), (14336, 4096), (5632, 8192), (8192, 8192), (28672, 8192), (4096, 14336), (2048, 5632)]:
        blk = 210 if q6 else 144
        wbytes = M * (K // 256) * blk
        copies = max(2, (320 << 20) // wbytes + 1)
        w = (rand_q6k if q6 else rand_q4k)(M, K, copies)
        x = torch.randn((1, K), device="cuda")
        dst = torch.empty((1, M), device="cuda")
        ws = torch.empty(max(q4k_b200.workspace_bytes(K, 1), 16), dtype=torch.uint8, device="cuda")
        f = q4k_b200.mul_mat_q6k if q6 else q4k_b200.mul_mat
        res = {}
        for name, lim in (("fused", 1 << 30), ("separate", 0)):
            if q6:
                q4k_b200.debug_set("q6k_fuse_limit", lim)
            else:
                lib.q4k_debug_set_fuse_limit(lim)
            us = graph_time_us(lambda i: f(w[i % copies], x, dst, ws), per_graph=max(copies, 20))
            res[name] = round(us, 2)
        lib.q4k_debug_set_fuse_limit(16384)
        q4k_b200.debug_set("q6k_fuse_limit", 32768)
        print(dict(type="q6_K" if q6 else "q4_K", M=M, K=K, **res, best_gbs=round((wbytes + K * 4 + M * 4) / min(res.values()) / 1e3, 1)), flush=True)
        del w
