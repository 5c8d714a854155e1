"""Q4_K N == 1 decode: the tensor-pipe kernel (gemv_q4k_mma.cu) against the lane-per-superblock bulk-copy kernel in the
same float order (gemv_q4k_tma.cu ORD 2) per shape — where is the crossover for q4k_debug_set_big_matrix_bytes?
CUDA-graph timed, weight copies rotated through > 2.5x L2.  Development tool."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from gemv_sweep import graph_time_us, rand_q4k  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shapes", default="2048x2048,5632x2048,2048x5632,8192x2048,4096x4096,14336x4096,4096x14336,8192x8192,28672x8192,8192x28672")
a = ap.parse_args()
lib = q4k_b200.load()
for shp in a.shapes.split(","):
    M, K = map(int, shp.split("x"))
    wbytes = M * (K // 256) * 144
    copies = max(2, (320 << 20) // wbytes + 1)
    w = rand_q4k(M, K, copies)
    x = torch.randn((1, K), device="cuda")
    dst = torch.empty((1, M), device="cuda")
    ws = torch.empty(q4k_b200.workspace_bytes(K, 1), dtype=torch.uint8, device="cuda")
    res = {}
    for name, thr in (("mma", 0), ("bulk", 1)):
        lib.q4k_debug_set_big_matrix_bytes(thr)
        us = graph_time_us(lambda i: q4k_b200.mul_mat(w[i % copies], x, dst, ws), per_graph=max(copies, 20))
        res[name] = (round(us, 2), round((wbytes + K * 4 + M * 4) / us / 1e3, 1))
    lib.q4k_debug_set_big_matrix_bytes(34 << 20)
    print(dict(M=M, K=K, MB=round(wbytes / 2**20, 1), mma_us=res["mma"][0], mma_gbs=res["mma"][1], bulk_us=res["bulk"][0],
               bulk_gbs=res["bulk"][1]), flush=True)
    del w
