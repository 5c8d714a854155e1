"""Per-launch cost of decode launches inside a stream-ordered chain (CUDA graph, PDL edges), per shape.
Each spec is M1+M2+...xK (a '+' list is one fused q4k_mul_mat_group launch).  Weight copies rotate through
>= 2.5x L2 so bytes come from HBM.  Development tool."""
import argparse
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import q4k_b200  # noqa: E402
from gemv_sweep import graph_time_us, rand_q4k  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--specs", default="4x2048,2048x2048,2048+256+256x2048,5632+5632x2048,2048x5632,5632x2048")
    ap.add_argument("--n", type=int, default=1)
    ap.add_argument("--tag", default="")
    ap.add_argument("--order", type=int, default=0, help="0 policy, 1 exact, 2 fast")
    ap.add_argument("--fuse-limit", type=int, default=-1, help="N*K floats up to which the quantizer is fused into the launch")
    a = ap.parse_args()
    lib = q4k_b200.load()
    lib.q4k_debug_set_gemv_order(a.order)
    if a.fuse_limit >= 0:
        lib.q4k_debug_set_fuse_limit(a.fuse_limit)
    N = a.n
    for spec in a.specs.split(","):
        ms, K = spec.split("x")
        Ms, K = [int(m) for m in ms.split("+")], int(K)
        wbytes = sum(Ms) * (K // 256) * 144
        copies = max(2, min(400, (320 << 20) // wbytes + 1))
        ws = [rand_q4k(M, K, copies) for M in Ms]
        x = torch.randn((N, K), device="cuda")
        dsts = [torch.empty((N, M), device="cuda") for M in Ms]
        wsb = torch.empty(max(q4k_b200.workspace_bytes(K, N), 16), dtype=torch.uint8, device="cuda")
        n = len(Ms)
        vpa, i64a = C.c_void_p * n, C.c_int64 * n

        def launch(i):
            st = torch.cuda.current_stream().cuda_stream
            if n == 1:
                rc = lib.q4k_mul_mat(ws[0][i % copies].data_ptr(), Ms[0], K, (K // 256) * 144, x.data_ptr(), N, K * 4,
                                     dsts[0].data_ptr(), Ms[0] * 4, wsb.data_ptr(), wsb.numel(), st)
            else:
                rc = lib.q4k_mul_mat_group(n, vpa(*[w[i % copies].data_ptr() for w in ws]), i64a(*Ms), K, (K // 256) * 144,
                                           vpa(*[x.data_ptr()] * n), N, K * 4, vpa(*[d.data_ptr() for d in dsts]), st)
            assert rc == 0, rc
        us = graph_time_us(launch, per_graph=max(copies, 50))
        print(dict(tag=a.tag, spec=spec, N=N, us=round(us, 3), gbs=round(wbytes / us / 1e3, 1), ideal_us=round(wbytes / 6550.7e3, 3)), flush=True)
        del ws
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
