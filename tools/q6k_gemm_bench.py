"""Q6_K prefill: the tcgen05 kind::f16 GEMM (csrc/gemm_q6k_tc.cu) against the decode kernel eight columns at a time, per shape.
Usage: python tools/q6k_gemm_bench.py [--iters 20]   (GPU).  Prints us per call and int TOPS (2*M*N*K / time)."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "ggml-neon-opt_b200"))
import q4k_b200  # noqa: E402

SHAPES = [(2048, 5632, 512), (256, 2048, 512), (32000, 2048, 512), (4096, 14336, 512), (2048, 5632, 64), (2048, 5632, 128),
          (32000, 2048, 33), (4096, 14336, 40), (2048, 5632, 8), (2048, 5632, 12), (2048, 5632, 16), (2048, 5632, 24), (2048, 5632, 32),
          (256, 2048, 12), (256, 2048, 16), (256, 2048, 24), (256, 2048, 32), (32000, 2048, 8), (32000, 2048, 12), (32000, 2048, 16), (32000, 2048, 24),
          (4096, 14336, 8), (4096, 14336, 16)]


def time_call(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--gemm-only", action="store_true")
    ap.add_argument("--first", type=int, default=len(SHAPES), help="only the first n shapes")
    args = ap.parse_args()
    L = q4k_b200.load()
    rng = np.random.default_rng(0)
    for M, K, N in SHAPES[:args.first]:
        nb = K // 256
        w = torch.from_numpy(rng.integers(0, 256, (M, nb, 210), dtype=np.uint8))
        w[:, :, 208:210] = torch.from_numpy(np.full((M, nb, 1), 0.001, np.float16).view(np.uint8).reshape(M, nb, 2))
        w = w.reshape(M, nb * 210).cuda()
        x = torch.randn(N, K, device="cuda")
        out = torch.empty(N, M, device="cuda")
        row = {}
        for name, gmin in ((("gemm", 1),) if args.gemm_only else (("gemm", 1), ("gemv_x8", 0))):
            q4k_b200.debug_set("q6k_gemm_min", gmin)
            nbytes = L.q6k_mul_mat_workspace_bytes(K, N)
            ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
            st = torch.cuda.current_stream().cuda_stream

            def call():
                rc = L.q6k_mul_mat(w.data_ptr(), M, K, w.stride(0), x.data_ptr(), N, K * 4, out.data_ptr(), M * 4, ws.data_ptr(), nbytes, st)
                assert rc == 0, rc
            row[name] = time_call(call, args.iters if name == "gemm" else max(3, args.iters // 4))
        q4k_b200.debug_set("q6k_gemm_min", 33)
        tops = 2.0 * M * N * K / row["gemm"] / 1e6
        tail = "" if args.gemm_only else f" | decode kernel x{(N + 7) // 8}: {row['gemv_x8']:.1f} us"
        print(f"{M}x{K} N={N}: gemm {row['gemm']:.1f} us = {tops:.0f} TOPS{tail}", flush=True)


if __name__ == "__main__":
    main()
