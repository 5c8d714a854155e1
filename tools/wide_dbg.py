"""Timing experiments on gemv_q4k_wide.cu: which part of a round costs what (knob wide_dbg: 1 = no activation copies, 2 = no
arithmetic, 4 = no weight copies; mma_stages caps the ring depths).  Development tool."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200
from gemv_sweep import graph_time_us, rand_q4k
q4k_b200.debug_set("wide_min", 1)
for M, K in [(28672, 8192), (4096, 14336)]:
    wbytes = M * (K // 256) * 144
    copies = max(2, (320 << 20) // wbytes + 1)
    w = rand_q4k(M, K, copies)
    for N in (4, 16, 32):
        x = torch.randn((N, K), device="cuda"); dst = torch.empty((N, M), device="cuda")
        ws = torch.empty(max(q4k_b200.workspace_bytes(K, N), q4k_b200.workspace_bytes(K, 9)), dtype=torch.uint8, device="cuda")
        res = {}
        for dbg in (0, 1, 2, 3, 4, 6, 7):
            q4k_b200.debug_set("wide_dbg", dbg)
            res[f"dbg{dbg}"] = round(graph_time_us(lambda i: q4k_b200.mul_mat(w[i % copies], x, dst, ws), per_graph=max(copies, 10), replays=5), 1)
        q4k_b200.debug_set("wide_dbg", 0)
        for st in (2, 3):
            q4k_b200.debug_set("mma_stages", st)
            res[f"stages{st}"] = round(graph_time_us(lambda i: q4k_b200.mul_mat(w[i % copies], x, dst, ws), per_graph=max(copies, 10), replays=5), 1)
        q4k_b200.debug_set("mma_stages", 0)
        print(json.dumps(dict(M=M, K=K, N=N, **res)), flush=True)
    del w
