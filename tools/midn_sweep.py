"""8 < N <= 128: the tcgen05 GEMM against the decode kernel run eight columns at a time (weights re-read from L2 by the
later passes), per shape.  Where should q4k_mul_mat switch?  Development tool."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from gemv_sweep import graph_time_us, rand_q4k  # noqa: E402

lib = q4k_b200.load()
lib.q4k_debug_set_gemm.argtypes = [__import__("ctypes").c_int]
for M, K in [(4096, 14336), (14336, 4096), (5632, 2048), (2048, 5632), (2048, 2048), (28672, 8192)]:
    wbytes = M * (K // 256) * 144
    copies = max(2, (320 << 20) // wbytes + 1)
    w = rand_q4k(M, K, copies)
    for N in (9, 16, 24, 32, 48, 64, 128):
        x = torch.randn((N, K), device="cuda")
        dst = torch.empty((N, M), device="cuda")
        ws = torch.empty(q4k_b200.workspace_bytes(K, N), dtype=torch.uint8, device="cuda")
        res = {}
        for name, on in (("gemm", 1), ("gemv_passes", 0)):
            lib.q4k_debug_set_gemm(on)
            res[name] = round(graph_time_us(lambda i: q4k_b200.mul_mat(w[i % copies], x, dst, ws), per_graph=max(copies, 10), replays=5), 1)
        lib.q4k_debug_set_gemm(1)
        print(dict(M=M, K=K, N=N, **res), flush=True)
    del w
