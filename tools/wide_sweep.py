"""8 < N <= 64: the batched tensor-pipe decode kernel (gemv_q4k_wide.cu) against what served these batches before it (two passes
of the decode kernel for N <= 12, the tcgen05 GEMM above), per shape, with the HBM fraction of the algorithmic bytes.
Development tool; output kept under profiles/."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from gemv_sweep import graph_time_us, rand_q4k  # noqa: E402

PEAK = 6459.0
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
shapes = [(4096, 14336), (14336, 4096), (5632, 2048), (2048, 5632), (28672, 8192)]
ns = [int(a) for a in sys.argv[1:]] or [4, 8, 9, 16, 24, 32, 64]
for M, K in shapes:
    wbytes = M * (K // 256) * 144
    copies = max(2, (320 << 20) // wbytes + 1)
    w = rand_q4k(M, K, copies)
    for N in ns:
        x = torch.randn((N, K), device="cuda")
        dst = torch.empty((N, M), device="cuda")
        ws = torch.empty(max(q4k_b200.workspace_bytes(K, N), q4k_b200.workspace_bytes(K, 9)), dtype=torch.uint8, device="cuda")
        res = {}
        for name, midn, wmin in (("wide", 64, 1), ("before", 0, 9)):      # N <= 8: "before" = the decode kernel (gemv_q4k_mma.cu)
            q4k_b200.debug_set("midn_max", midn)
            q4k_b200.debug_set("wide_min", wmin)
            res[name] = round(graph_time_us(lambda i: q4k_b200.mul_mat(w[i % copies], x, dst, ws), per_graph=max(copies, 10), replays=5), 1)
        q4k_b200.debug_set("midn_max", 64)
        q4k_b200.debug_set("wide_min", 2)
        algo = wbytes + N * K * 4 + M * N * 4
        gbs = algo / res["wide"] / 1e3
        print(json.dumps(dict(M=M, K=K, N=N, us_wide=res["wide"], us_before=res["before"], gbs=round(gbs, 1), frac_measured=round(gbs / PEAK, 3))), flush=True)
    del w
