"""GEMV sweep on one GPU: achieved GB/s of algorithmic bytes per shape / batch / float-order policy.
Weights rotate through >= 2x L2 of distinct copies so the numbers are HBM, not L2.  Development tool."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
import q4k_b200  # noqa: E402


def rand_q4k(M, K, copies):
    nb = K // 256
    w = torch.randint(0, 256, (copies, M, nb, 144), dtype=torch.uint8, device="cuda")
    w[..., 0] = 0x00; w[..., 1] = 0x2c; w[..., 2] = 0x00; w[..., 3] = 0x28      # d = 2^-4, dmin = 2^-5 (fp16)
    return w.view(copies, M, nb * 144)


def time_ms(fn, iters):
    for _ in range(5):
        fn(0)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(True), torch.cuda.Event(True)
    s.record()
    for i in range(iters):
        fn(i)
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def main():
    shapes = [(5632, 2048), (2048, 5632), (4096, 14336), (14336, 4096), (28672, 8192), (8192, 28672)]
    out = []
    lib = q4k_b200.load()
    for M, K in shapes:
        wbytes = M * (K // 256) * 144
        copies = max(2, (320 << 20) // wbytes + 1)
        w = rand_q4k(M, K, copies)
        for N in (1, 2, 3, 4, 8):
            x = torch.randn((N, K), device="cuda")
            xq = q4k_b200.quantize_q8_K(x)
            dst = torch.empty((N, M), device="cuda")
            for order in ("exact", "fast"):
                exact = order == "exact"
                st = torch.cuda.current_stream().cuda_stream

                def run(i, exact=exact):
                    os.environ  # noqa
                    rc = lib.q4k_mul_mat_q8(w[i % copies].data_ptr(), M, K, (K // 256) * 144, xq.data_ptr(), N, xq.stride(0),
                                            dst.data_ptr(), M * 4, st)
                    assert rc == 0
                os.environ["Q4K_GEMV_ORDER"] = order
                # the policy env var is read once per process; use the per-call override symbol instead
                lib.q4k_debug_set_gemv_order(1 if exact else 2)
                ms = time_ms(run, 100)
                algo = wbytes + N * K * 4 + M * N * 4
                out.append(dict(M=M, K=K, N=N, order=order, us=round(ms * 1e3, 2), gbs=round(algo / ms / 1e6, 1)))
                print(out[-1], flush=True)
        del w
        torch.cuda.empty_cache()
    lib.q4k_debug_set_gemv_order(0)
    # quantizer
    for N, K in ((1, 2048), (1, 14336), (512, 2048), (512, 14336)):
        x = torch.randn((N, K), device="cuda")
        y = torch.empty((N, K // 256 * 292), dtype=torch.uint8, device="cuda")
        ms = time_ms(lambda i: q4k_b200.quantize_q8_K(x, y), 100)
        print(dict(quantize=(N, K), us=round(ms * 1e3, 2), gbs=round((N * K * 4 + y.numel()) / ms / 1e6, 1)), flush=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gemv_sweep.json"), "w"))


if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    main()
