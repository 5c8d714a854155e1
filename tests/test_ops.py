"""The cheap F32 ops between the mul_mats of a decoder layer (include/q4k_b200.h: q4k_op_*):
RMS_NORM (+ fused residual ADD / weight MUL), ADD, MUL, GLU/SWIGLU, GET_ROWS, ROPE.

CPU: oracle/port.py's restatements against vectors made by the reference's own CPU backend (tests/golden/ops_ref_vectors.npz,
tests/golden/make_golden.py) and, where oracle/_ref is present, against the reference itself.
GPU: the kernels through the C ABI against the same vectors and the restatements, and the plugin under the reference's own
test-backend-ops (which compares with ggml_backend_cpu, NMSE <= 1e-7 for these ops).

Tolerances.  ADD / MUL: bit-exact (one IEEE operation).  RMS_NORM: <= 1 ulp per element (2^-23 relative) — the double sum runs
in a tree on the GPU and in element order on the CPU, which can move the float mean by one rounding at most.  SWIGLU: 4e-6
relative (+1e-30 absolute) — the reference's AVX2 silu uses a polynomial expf.  ROPE: 1e-6 of the row maximum (a few ulp of the
pair magnitude: the angle is built with the reference's own float products, cosf / sinf differ by an ulp or two)."""
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import port, ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ops_ref_vectors.npz"))
PLUGIN = os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libggml-q4k-b200.so")
TBO = os.path.join(ROOT, "oracle", "_ref", "test-backend-ops")

ULP = 2.0 ** -23
SILU_RTOL = 4e-6


def rel_err(got, want):
    return np.abs(got.astype(np.float64) - want) / np.maximum(np.abs(want.astype(np.float64)), 1e-30)


# ---------------------------------------------------------------------------------------------------- CPU
def test_port_matches_reference_vectors():
    s = (G["x"] + G["res"]).astype(np.float32)
    assert np.array_equal(s.view(np.uint32), G["y_sum"].view(np.uint32))
    n = (port.rms_norm(s, float(G["eps"])) * G["w"]).astype(np.float32)
    assert np.array_equal(n.view(np.uint32), G["y_norm"].view(np.uint32))          # same summation order as the reference: same bits
    g = port.swiglu(G["g"], G["u"])
    assert rel_err(g, G["y_glu"]).max() <= SILU_RTOL


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
def test_port_matches_reference_run_here():
    rng = np.random.default_rng(9)
    rows, ne0 = 3, 5632
    x, r = rng.standard_normal((rows, ne0)).astype(np.float32), rng.standard_normal((rows, ne0)).astype(np.float32)
    w = rng.uniform(0.5, 1.5, ne0).astype(np.float32)
    g, u = (rng.standard_normal((rows, ne0)) * 3).astype(np.float32), rng.standard_normal((rows, ne0)).astype(np.float32)
    ys, yn, yg = ref.cheap_ops(x, r, w, g, u, 1e-6)
    assert np.array_equal((x + r).astype(np.float32).view(np.uint32), ys.view(np.uint32))
    assert np.array_equal((port.rms_norm(ys, 1e-6) * w).astype(np.float32).view(np.uint32), yn.view(np.uint32))
    assert rel_err(port.swiglu(g, u), yg).max() <= SILU_RTOL


def rope_close(got, want, x):
    """a few ulp of the pair magnitude: cosf / sinf are within 1-2 ulp, the rotation may or may not contract to FMAs"""
    rowmax = np.abs(x).max(axis=-1, keepdims=True)
    return np.all(np.abs(got.astype(np.float64) - want) <= 1e-6 * rowmax)


def test_port_rope_matches_reference_vectors():
    x, pos = G["rope_x"], G["rope_pos"]
    assert port.rope_theta_scale(10000.0, 64) == G["rope_theta_scale"]
    assert rope_close(port.rope(x, pos, 64), G["rope_y_normal"], x)
    assert rope_close(port.rope(x, pos, 64, neox=True), G["rope_y_neox"], x)
    assert np.array_equal(G["rope_y_normal"][0], x[0])                       # position 0: identity


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
def test_port_rope_partial_dims_and_scaling_against_reference_run_here():
    rng = np.random.default_rng(4)
    x, pos = rng.standard_normal((3, 2, 128)).astype(np.float32), np.array([5, 900, 4000], np.int32)
    for mode in (0, 2):
        want = ref.rope(x, pos, 64, mode, freq_base=500000.0, freq_scale=0.5)        # n_dims < head_dim: the tail is copied
        assert rope_close(port.rope(x, pos, 64, neox=mode == 2, freq_base=500000.0, freq_scale=0.5), want, x)
        assert np.array_equal(want[..., 64:], x[..., 64:])


# ---------------------------------------------------------------------------------------------------- GPU
def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
def test_gpu_ops_match_reference_vectors(q4k):
    x, r, w = dev(G["x"]), dev(G["res"]), dev(G["w"])
    eps = float(G["eps"])
    # separate launches
    s = q4k.op_binary("add", x, r)
    assert np.array_equal(s.cpu().numpy().view(np.uint32), G["y_sum"].view(np.uint32))
    n = q4k.op_binary("mul", q4k.op_rms_norm(s, eps), w)
    assert rel_err(n.cpu().numpy(), G["y_norm"]).max() <= 2 * ULP
    # one launch: ADD -> RMS_NORM -> MUL
    y, s2 = q4k.op_rms_norm(x, eps, weight=w, residual=r, want_sum=True)
    assert np.array_equal(s2.cpu().numpy().view(np.uint32), G["y_sum"].view(np.uint32))
    assert np.array_equal(y.cpu().numpy().view(np.uint32), n.cpu().numpy().view(np.uint32))      # fused == separate, bit for bit
    y_nosum = q4k.op_rms_norm(x, eps, weight=w, residual=r)
    assert np.array_equal(y_nosum.cpu().numpy().view(np.uint32), y.cpu().numpy().view(np.uint32))
    g = q4k.op_swiglu(dev(G["g"]), dev(G["u"]))
    assert rel_err(g.cpu().numpy(), G["y_glu"]).max() <= SILU_RTOL


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 2048), (512, 2048), (3, 7, 100), (2, 3, 4, 33), (1, 1)])
def test_gpu_rms_norm_shapes(q4k, shape):
    import torch
    rng = np.random.default_rng(sum(shape))
    x = rng.standard_normal(shape).astype(np.float32)
    got = q4k.op_rms_norm(dev(x), 1e-5).cpu().numpy()
    assert rel_err(got, port.rms_norm(x, 1e-5)).max() <= 2 * ULP
    # a strided view: rows of a wider buffer (what a ggml view hands over)
    wide = torch.from_numpy(rng.standard_normal(shape[:-1] + (shape[-1] + 12,)).astype(np.float32)).cuda()
    view = wide[..., 4:4 + shape[-1]]
    got = q4k.op_rms_norm(view, 1e-6).cpu().numpy()
    assert rel_err(got, port.rms_norm(view.cpu().numpy(), 1e-6)).max() <= 2 * ULP


@pytest.mark.gpu
@pytest.mark.parametrize("op", ["add", "mul"])
@pytest.mark.parametrize("a_shape,b_shape", [((4, 2048), (4, 2048)), ((4, 2048), (1, 2048)), ((4, 2048), (2048,)), ((2, 3, 4, 10), (1, 3, 1, 10)),
                                             ((2, 3, 4, 10), (2, 1, 2, 5)), ((5, 64), (5, 1)), ((512, 5632), (5632,))])
def test_gpu_binary_broadcast_bit_exact(q4k, op, a_shape, b_shape):
    rng = np.random.default_rng(len(a_shape) * 100 + len(b_shape))
    a, b = rng.standard_normal(a_shape).astype(np.float32), rng.standard_normal(b_shape).astype(np.float32)
    got = q4k.op_binary(op, dev(a), dev(b)).cpu().numpy()
    b4 = b.reshape((1,) * (len(a_shape) - len(b_shape)) + b_shape)
    reps = tuple(sa // sb for sa, sb in zip(a_shape, b4.shape))
    bt = np.tile(b4, reps)                                         # ggml repeat semantics: index modulo the extent
    want = (a + bt if op == "add" else a * bt).astype(np.float32)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("rows,nc", [(1, 5632), (8, 5632), (512, 5632), (3, 17)])
def test_gpu_swiglu_and_get_rows(q4k, rows, nc):
    import torch
    rng = np.random.default_rng(rows + nc)
    g, u = (rng.standard_normal((rows, nc)) * 5).astype(np.float32), rng.standard_normal((rows, nc)).astype(np.float32)
    got = q4k.op_swiglu(dev(g), dev(u)).cpu().numpy()
    assert rel_err(got, port.swiglu(g, u)).max() <= SILU_RTOL
    ids = rng.integers(0, rows, 2 * rows + 1).astype(np.int32)
    picked = q4k.op_get_rows(dev(g), torch.from_numpy(ids).cuda()).cpu().numpy()
    assert np.array_equal(picked.view(np.uint32), g[ids].view(np.uint32))


@pytest.mark.gpu
def test_gpu_rope_matches_reference_vectors(q4k):
    import torch
    x, pos = G["rope_x"], G["rope_pos"]
    ts = float(G["rope_theta_scale"])
    for neox, key in ((False, "rope_y_normal"), (True, "rope_y_neox")):
        got = q4k.op_rope(dev(x), torch.from_numpy(pos).cuda(), 64, ts, neox=neox).cpu().numpy()
        assert rope_close(got, G[key], x)
    # partial rotation, frequency scaling, a strided (view) source; checked against the restatement
    rng = np.random.default_rng(8)
    wide = rng.standard_normal((5, 3, 160)).astype(np.float32)
    view = torch.from_numpy(wide).cuda()[..., 16:144]
    p2 = np.array([3, 77, 1000, 4095, 8191], np.int32)
    ts2 = float(port.rope_theta_scale(500000.0, 64))
    for neox in (False, True):
        got = q4k.op_rope(view, torch.from_numpy(p2).cuda(), 64, ts2, neox=neox, freq_scale=0.5).cpu().numpy()
        want = port.rope(wide[..., 16:144], p2, 64, neox=neox, freq_base=500000.0, freq_scale=0.5)
        assert rope_close(got, want, wide[..., 16:144])
        assert np.array_equal(got[..., 64:], wide[..., 80:144])


@pytest.mark.gpu
def test_ops_reject_bad_arguments(q4k):
    import ctypes as C
    lib = q4k.load()
    ne = (C.c_int64 * 4)(8, 1, 1, 1)
    nb = (C.c_int64 * 3)(32, 32, 32)
    assert lib.q4k_op_rms_norm(None, nb, None, None, None, None, nb, None, None, ne, 1e-5, None) == -2
    assert lib.q4k_op_binary(7, None, ne, nb, None, ne, nb, None, nb, None) == -5
    bad = (C.c_int64 * 4)(3, 1, 1, 1)
    x = dev(np.zeros(8, np.float32))
    assert lib.q4k_op_binary(0, x.data_ptr(), ne, nb, x.data_ptr(), bad, nb, x.data_ptr(), nb, None) == -1     # 8 % 3: cannot repeat


@pytest.mark.gpu
@pytest.mark.skipif(not (ref.available() and os.path.exists(PLUGIN) and os.path.exists(TBO)), reason="oracle/_ref or the plugin not built")
@pytest.mark.parametrize("op,min_ok", [("RMS_NORM", 4), ("ADD", 8), ("MUL", 8), ("SWIGLU", 2), ("GET_ROWS", 2), ("ROPE", 8)])
def test_reference_test_backend_ops_cheap_ops(op, min_ok):
    """The reference's own harness (L/tests/test-backend-ops.cpp) on the plugin: every case the plugin claims is compared with
    ggml_backend_cpu; cases outside the supports_op contract (other types, permuted operands) report 'not supported'."""
    env = dict(os.environ, GGML_BACKEND_PATH=PLUGIN)
    p = subprocess.run([TBO, "test", "-o", op, "-b", "Q4K_B200_0"], env=env, capture_output=True, text=True, timeout=900)
    out = re.sub(r"\x1b\[[0-9;]*m", "", p.stdout + p.stderr)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    open(os.path.join(ROOT, "gpurun_out", f"test_backend_ops_{op.lower()}.log"), "w").write(out)
    fails = [l for l in out.splitlines() if "FAIL" in l]
    oks = [l for l in out.splitlines() if l.rstrip().endswith("OK") and op in l]
    assert p.returncode == 0 and not fails, (fails[:5], out[-1500:])
    assert len(oks) >= min_ok, out[-1500:]
