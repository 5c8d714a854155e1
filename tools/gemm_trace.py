"""Per-role clock64 stamps of CTA (0,0) of the prefill GEMM (make -C ggml-neon-opt_b200/csrc trace).  Development tool."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["Q4K_B200_LIB"] = os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libq4k_b200_trace.so")
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from gemv_sweep import rand_q4k  # noqa: E402

M, K, N = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "4096x14336x512").split("x"))
w = rand_q4k(M, K, 1)[0]
x = torch.randn((N, K), device="cuda")
out = torch.empty((N, M), device="cuda")
ws = torch.empty(q4k_b200.workspace_bytes(K, N), dtype=torch.uint8, device="cuda")
for _ in range(3):
    q4k_b200.mul_mat(w, x, out, ws)
torch.cuda.synchronize()
L = q4k_b200.load()
buf = np.zeros((16, 256), dtype=np.int64)
L.q4k_debug_gemm_trace.argtypes = [ctypes.c_void_p]
assert L.q4k_debug_gemm_trace(buf.ctypes.data) == 0
nsteps = min(2 * K // 256, 256)
nb = nsteps // 2
t0 = buf[0, 0]
names = {0: "P:loads landed", 1: "P:emptyA ok", 2: "P:full arrive", 3: "P:emptyB ok", 4: "M:top", 5: "M:full ok", 6: "M:tempty0 ok",
         7: "M:tempty1 ok", 8: "E0:tfull ok", 9: "E0:arrive", 10: "E1:tfull ok", 11: "E1:arrive"}
print("total cycles (producer step 0 -> last full arrive):", buf[2, nsteps - 1] - t0, " per step:", (buf[2, nsteps - 1] - t0) / nsteps)
lo, hi = nsteps // 2, nsteps // 2 + 8
print("steady-state window, cycles relative to P:loads landed of step", lo)
base = buf[0, lo]
for s in range(lo, hi):
    kb = s // 2
    row = [f"s{s:3d}"] + [f"{names[i]}={buf[i, s] - base:6d}" for i in (0, 1, 2, 3, 4, 5)]
    if s % 2 == 0:
        row += [f"{names[i]}={buf[i, s] - base:6d}" for i in (6, 7)]
    else:
        row += [f"{names[i]}={buf[i, kb] - base:6d}" for i in (8, 9, 10, 11)]
    print("  ".join(row))
# averages over the steady state
S = np.arange(8, nsteps - 8)
def avg(a): return float(np.mean(a))
print("avg per step: unpack (emptyA ok -> full arrive) %.0f   wait emptyA %.0f   wait emptyB %.0f   loads+bar %.0f" % (
    avg(buf[2, S] - buf[1, S]), avg(buf[1, S] - buf[0, S]), avg(buf[3, S] - buf[2, S]), avg(buf[0, S + 1] - buf[3, S])))
print("avg MMA thread: wait full %.0f   step period %.0f" % (avg(buf[5, S] - buf[4, S]), avg(buf[4, S + 1] - buf[4, S])))
Kb = np.arange(4, nb - 4)
print("avg MMA thread on even steps: full ok -> tempty0 ok %.0f   tempty0 ok -> tempty1 ok %.0f" % (
    avg(buf[6, 2 * Kb] - buf[5, 2 * Kb]), avg(buf[7, 2 * Kb] - buf[6, 2 * Kb])))
print("avg epilogue: E0 drain (tfull ok -> arrive) %.0f  E1 drain %.0f   E0 wait (arrive kb-1 -> tfull ok kb) %.0f  E1 wait %.0f" % (
    avg(buf[9, Kb] - buf[8, Kb]), avg(buf[11, Kb] - buf[10, Kb]), avg(buf[8, Kb] - buf[9, Kb - 1]), avg(buf[10, Kb] - buf[11, Kb - 1])))
print("avg epilogue: h0 math after arrive %.0f   h1 math after arrive %.0f" % (avg(buf[12, Kb] - buf[9, Kb]), avg(buf[14, Kb] - buf[11, Kb])))
print("avg: tfull0 ok (epilogue) - M issue of odd step top %.0f ; tempty0 ok (MMA) - E0 arrive %.0f" % (
    avg(buf[8, Kb] - buf[5, 2 * Kb + 1]), avg(buf[6, 2 * Kb + 2] - buf[9, Kb])))
