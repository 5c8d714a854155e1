"""The ggml backend plugin (ggml-neon-opt_b200/lib/libggml-q4k-b200.so) driven by the UNMODIFIED reference:
loaded through ggml_backend_load / GGML_BACKEND_PATH (L/ggml/src/ggml-backend-reg.cpp:247-291, 603-607), run by
the reference's own graph API and by its own per-op harness test-backend-ops (L/tests/test-backend-ops.cpp),
which compares every case with the reference CPU backend (NMSE <= 5e-4 for MUL_MAT, :3300-3302)."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import make_acts, make_weights
from oracle import port, ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libggml-q4k-b200.so")
TBO = os.path.join(ROOT, "oracle", "_ref", "test-backend-ops")

needs_ref = pytest.mark.skipif(not (ref.available() and os.path.exists(PLUGIN)), reason="oracle/_ref or the plugin not built")


@needs_ref
def test_plugin_exports_dl_entry_points():
    out = subprocess.check_output(["nm", "-D", "--defined-only", PLUGIN], text=True)
    assert " T ggml_backend_init" in out and " T ggml_backend_score" in out


@needs_ref
def test_plugin_scores_zero_without_sm100_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    w, x = make_weights(4, 256), make_acts(1, 256, adversarial=False)
    with pytest.raises(RuntimeError, match="rc=-10"):      # ggml_backend_load refuses a backend whose score is 0
        ref.plugin_mul_mat(PLUGIN, w, x)


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("M,K,N", [(5632, 2048, 1), (300, 1024, 3), (128, 512, 16)])
def test_plugin_graph_matches_reference_cpu_backend(M, K, N):
    import ctypes
    w, x = make_weights(M, K), make_acts(N, K)
    cpu, _, _ = ref.mul_mat(w, x, n_threads=4)
    f64 = port.mul_mat_f64(w, x)
    rms = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
    got, _, _ = ref.plugin_mul_mat(PLUGIN, w, x)                              # default float order: within 1e-5
    assert np.all(np.abs(got - cpu) <= 1e-5 * np.maximum(np.abs(cpu), rms))
    if N <= 8:                                                               # AVX2-order decode kernel: bit-identical
        lib = ctypes.CDLL(os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libq4k_b200.so"))   # the instance the plugin links
        lib.q4k_debug_set_gemv_order(1)
        try:
            got, _, _ = ref.plugin_mul_mat(PLUGIN, w, x)
        finally:
            lib.q4k_debug_set_gemv_order(0)
        if N <= 2:
            assert np.array_equal(got.view(np.uint32), cpu.view(np.uint32))
        else:
            assert np.all(np.abs(got - cpu) <= 1e-5 * np.maximum(np.abs(cpu), rms))


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("N", [1, 4, 16])
def test_plugin_fuses_mul_mats_on_one_activation_tensor(N):
    """q / k / v of a layer: three ggml_mul_mat nodes on one `cur`.  graph_compute sends them out as one launch
    (q4k_mul_mat_group) for N <= 8 and one by one above; either way every result equals the single-node graph's bits."""
    import ctypes
    K = 1024
    ws = [make_weights(M, K, seed=21 + i) for i, M in enumerate((512, 128, 96))]
    x = make_acts(N, K)
    lib = ctypes.CDLL(os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libq4k_b200.so"))
    lib.q4k_kernel_launches.restype = ctypes.c_int64
    n0 = lib.q4k_kernel_launches()
    outs = ref.plugin_mul_mat_shared(PLUGIN, ws, x)
    launches = lib.q4k_kernel_launches() - n0
    if N <= 8:
        assert launches == 1, launches
    for w, o in zip(ws, outs):
        alone, _, _ = ref.plugin_mul_mat(PLUGIN, w, x)
        assert np.array_equal(o.view(np.uint32), alone.view(np.uint32))
        cpu, _, _ = ref.mul_mat(w, x, n_threads=4)
        f64 = port.mul_mat_f64(w, x)
        rms = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
        assert np.all(np.abs(o - cpu) <= 1e-5 * np.maximum(np.abs(cpu), rms))


@needs_ref
@pytest.mark.gpu
def test_reference_test_backend_ops_mul_mat():
    env = dict(os.environ, GGML_BACKEND_PATH=PLUGIN)
    p = subprocess.run([TBO, "test", "-o", "MUL_MAT", "-b", "Q4K_B200_0"], env=env, capture_output=True, text=True, timeout=900)
    out = p.stdout + p.stderr
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    open(os.path.join(ROOT, "gpurun_out", "test_backend_ops_mul_mat.log"), "w").write(out)
    clean = re.sub(r"\x1b\[[0-9;]*m", "", out)
    q4k_ok = [l for l in clean.splitlines() if "type_a=q4_K" in l and l.rstrip().endswith("OK")]
    fails = [l for l in clean.splitlines() if "FAIL" in l]
    assert p.returncode == 0, clean[-2000:]
    assert not fails, fails[:5]
    assert len(q4k_ok) >= 9, clean[-2000:]          # at least the m=16, n=1..9, k=256 sweep ran on the GPU
    q6k_ok = [l for l in clean.splitlines() if "type_a=q6_K" in l and l.rstrip().endswith("OK")]
    assert len(q6k_ok) >= 9, clean[-2000:]          # the plugin claims MUL_MAT(Q6_K) too
    assert re.search(r"Backend Q4K_B200: OK", clean)
