"""One Q6_K prefill shape through q6k_mul_mat, a few calls (for ncu).  Usage: python tools/q6k_gemm_one.py M K N [iters]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "ggml-neon-opt_b200"))
import q4k_b200  # noqa: E402

M, K, N = (int(a) for a in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
L = q4k_b200.load()
nb = K // 256
rng = np.random.default_rng(0)
w = torch.from_numpy(rng.integers(0, 256, (M, nb, 210), dtype=np.uint8))
w[:, :, 208:210] = torch.from_numpy(np.full((M, nb, 1), 0.001, np.float16).view(np.uint8).reshape(M, nb, 2))
w = w.reshape(M, nb * 210).cuda()
x = torch.randn(N, K, device="cuda")
out = torch.empty(N, M, device="cuda")
nbytes = L.q6k_mul_mat_workspace_bytes(K, N)
ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
for _ in range(iters):
    assert L.q6k_mul_mat(w.data_ptr(), M, K, w.stride(0), x.data_ptr(), N, K * 4, out.data_ptr(), M * 4, ws.data_ptr(), nbytes,
                         torch.cuda.current_stream().cuda_stream) == 0
torch.cuda.synchronize()
print("ok", float(out.abs().mean()))
