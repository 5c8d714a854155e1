"""Generates tests/golden/q4k_ref_vectors.npz by running the UNMODIFIED reference (oracle/_ref, built by
oracle/Makefile from /root/reference) on seeded inputs.  Run in the build container:
    python tests/golden/make_golden.py
The reference has no golden vectors of its own for this path; these pin the oracle and
the CUDA path where /root/reference is not available (the GPU box)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402

rng = np.random.default_rng(20261104)
M, K, N = 12, 1024, 4
w = rng.uniform(-1, 1, (M, K)).astype(np.float32)
w[3, :256] = 0.0                       # an all-zero weight superblock
w[4, 256:288] = 0.25                   # a constant sub-block (max == min branch after min clamp)
w[5] *= 1e-3
x = rng.standard_normal((N, K)).astype(np.float32)
x[0, :256] = 0.0
x[1, 256] = -9.0
x[1, 257] = 9.0
x[2, 512:768] = np.float32(1e-39)
x[3] *= 100.0
# the test-quantize-fns synthetic row (L/tests/test-quantize-fns.cpp:24-28): 0.1 + 2cos(i + offset)
x_cos = (0.1 + 2.0 * np.cos(np.arange(K, dtype=np.float32) + np.float32(1.0))).astype(np.float32)[None, :]
x = np.concatenate([x, x_cos], axis=0)

wq = ref.quantize_q4_K(w)
xq = ref.quantize_q8_K(x)
y_avx2, _, _ = ref.mul_mat(wq, x, n_threads=2)
y_direct = ref.mul_mat_direct(wq, x)
assert np.array_equal(y_avx2.view(np.uint32), y_direct.view(np.uint32))
y_generic = np.array([[ref.vec_dot(wq[m], xq[n], generic=True) for m in range(M)] for n in range(x.shape[0])], np.float32)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "q4k_ref_vectors.npz"),
                    w=w, x=x, wq=wq, xq=xq, y_avx2=y_avx2, y_generic=y_generic, w_deq=ref.dequantize_q4_K(wq),
                    note=np.array("llama.cpp a3cb0474, ggml-base no -m flags, ggml-cpu -mavx2 -mfma -mf16c -mbmi2, gcc 13.3"))
print("wrote", wq.shape, xq.shape, y_avx2.shape)

# ---- Q6_K: same inputs through quantize_row_q6_K_ref / ggml_vec_dot_q6_K_q8_K
w6 = ref.quantize_q6_K(w)
y6_avx2 = np.array([[ref.vec_dot_q6(w6[m], xq[n]) for m in range(M)] for n in range(x.shape[0])], np.float32)
y6_generic = np.array([[ref.vec_dot_q6(w6[m], xq[n], generic=True) for m in range(M)] for n in range(x.shape[0])], np.float32)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "q6k_ref_vectors.npz"),
                    w=w, x=x, wq=w6, xq=xq, y_avx2=y6_avx2, y_generic=y6_generic, w_deq=ref.dequantize_q6_K(w6),
                    note=np.array("llama.cpp a3cb0474, ggml-base no -m flags, ggml-cpu -mavx2 -mfma -mf16c -mbmi2, gcc 13.3"))
print("wrote q6", w6.shape, y6_avx2.shape)

from oracle.port import rope_theta_scale as port_theta_scale  # noqa: E402

# ---- the cheap F32 ops of a decoder layer: ADD -> RMS_NORM -> MUL and SWIGLU on ggml_backend_cpu
r2 = np.random.default_rng(20261105)
rows, ne0 = 5, 2048
ox, ores = r2.standard_normal((rows, ne0)).astype(np.float32), r2.standard_normal((rows, ne0)).astype(np.float32)
ox[1] *= 1e-4; ores[1] *= 1e-4          # tiny row: eps dominates the mean
ox[2] *= 300.0                          # large row
ow = r2.uniform(0.5, 1.5, ne0).astype(np.float32)
og, ou = (r2.standard_normal((rows, ne0)) * 4).astype(np.float32), r2.standard_normal((rows, ne0)).astype(np.float32)
og[3, :64] = np.linspace(-90, 90, 64, dtype=np.float32)       # silu tails: expf overflow / underflow side
ys, yn, yg = ref.cheap_ops(ox, ores, ow, og, ou, 1e-5)
# ROPE as the llama graph calls it (head_dim 64, freq_base 10000, NORMAL mode) and the NEOX pairing; positions up to the context end
rx = r2.standard_normal((6, 4, 64)).astype(np.float32)
rpos = np.array([0, 1, 17, 127, 511, 2047], np.int32)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "ops_ref_vectors.npz"), x=ox, res=ores, w=ow, g=og, u=ou, eps=np.float32(1e-5),
                    y_sum=ys, y_norm=yn, y_glu=yg, rope_x=rx, rope_pos=rpos, rope_theta_scale=port_theta_scale(10000.0, 64),
                    rope_y_normal=ref.rope(rx, rpos, 64, 0), rope_y_neox=ref.rope(rx, rpos, 64, 2),
                    note=np.array("llama.cpp a3cb0474 ggml_backend_cpu: add, rms_norm, mul, swiglu_split; ggml-cpu -mavx2 -mfma -mf16c -mbmi2"))
print("wrote ops", ys.shape)
