"""The plugin's split (row-sharded) buffer type — SURVEY.md §8e inside the C / C++ boundary — driven by the UNMODIFIED reference the
way llama.cpp's --split-mode row drives ggml-cuda's (L/src/llama-model.cpp:375-401, L/ggml/src/ggml-cuda/ggml-cuda.cu:795-1050):
weights in the buffer type returned by the registry's "ggml_backend_split_buffer_type" proc, one ggml_mul_mat node on the main
device's backend.  Results must equal the unsharded plugin result BIT FOR BIT (a row's bits do not depend on where it sits) and
the reference CPU backend within 1e-5.  With one GPU the split has one part (still the whole mechanism: fake base address, shards
in ->extra, scatter / gather, event fork / join); with more, every device takes a row range."""
import os

import numpy as np
import pytest

from conftest import make_acts, make_weights
from oracle import port, ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libggml-q4k-b200.so")

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (ref.available() and os.path.exists(PLUGIN)), reason="oracle/_ref or the plugin not built")]


@pytest.mark.parametrize("M,K,N", [(300, 1024, 1), (5632, 2048, 1), (100, 2048, 4), (257, 512, 16), (4096, 14336, 1), (160, 1024, 40)])
def test_split_buffer_mul_mat_equals_the_unsharded_plugin(M, K, N):
    w, x = make_weights(M, K, seed=M + N), make_acts(N, K, seed=K + N)
    import torch
    y, wb, nd = ref.plugin_split_mul_mat(PLUGIN, w, x)
    assert nd == torch.cuda.device_count()                                  # one plugin device per sm_100 GPU: the split covers them all
    assert np.array_equal(wb, w)                                            # get_tensor gathers what set_tensor scattered
    whole, _, _ = ref.plugin_mul_mat(PLUGIN, w, x)
    if N <= 32:                              # (the GEMM's split-K choice depends on the row count: tolerance there)
        assert np.array_equal(y.view(np.uint32), whole.view(np.uint32))
    cpu, _, _ = ref.mul_mat(w, x, n_threads=4)
    f64 = port.mul_mat_f64(w, x)
    rms = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
    assert np.all(np.abs(y - cpu) <= 1e-5 * np.maximum(np.abs(cpu), rms))
    assert np.all(np.abs(y - whole) <= 1e-5 * np.maximum(np.abs(whole), rms))


def test_split_buffer_uneven_shares_and_q6k():
    import torch
    nd = torch.cuda.device_count()
    M, K, N = 1000, 768, 2
    w6 = port.random_q6_K(np.random.default_rng(5), M, K)
    x = make_acts(N, K, seed=9)
    shares = [3.0, 1.0] + [0.0] * 14 if nd >= 2 else None                   # 3 : 1 over the first two devices, nothing for the rest
    y, wb, _ = ref.plugin_split_mul_mat(PLUGIN, w6, x, q6k=True, tensor_split=shares)
    assert np.array_equal(wb, w6)
    want = port.mul_mat_q6(w6, x, port.ORDER_AVX2)
    f64 = port.mul_mat_q6_f64(w6, x)
    rms = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
    assert np.all(np.abs(y - want) <= 1e-5 * np.maximum(np.abs(want), rms))


def test_split_ranges_c_abi(q4k):
    """q4k_split_create: contiguous, granule-aligned, covering row ranges; equal shares by default"""
    import ctypes as C
    import torch
    lib = q4k.load()
    lib.q4k_split_create.restype = C.c_void_p
    lib.q4k_split_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_float), C.c_int64, C.c_int64, C.c_int]
    lib.q4k_split_range.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.q4k_split_destroy.argtypes = [C.c_void_p]
    lib.q4k_split_parts.argtypes = [C.c_void_p]
    nd = torch.cuda.device_count()
    devs = (C.c_int * 4)(*[i % nd for i in range(4)])                        # four parts, cycling over the devices there are
    fr = (C.c_float * 4)(1.0, 2.0, 0.0, 1.0)
    for fractions, M in ((None, 1000), (fr, 1000), (None, 7), (fr, 16)):
        s = lib.q4k_split_create(devs, 4, fractions, M, 144, 16)
        assert s and lib.q4k_split_parts(s) == 4
        prev = 0
        for i in range(4):
            d, lo, hi = C.c_int(), C.c_int64(), C.c_int64()
            assert lib.q4k_split_range(s, i, C.byref(d), C.byref(lo), C.byref(hi)) == 0
            assert lo.value == prev and hi.value >= lo.value and (hi.value % 16 == 0 or i == 3)
            prev = hi.value
        assert prev == M
        lib.q4k_split_destroy(s)
