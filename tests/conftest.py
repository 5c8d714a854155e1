import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref (the reference compiled by oracle/Makefile)")


def make_weights(M, K, seed=1234):
    """U(-1,1) float32 -> Q4_K bytes with the oracle's restatement of quantize_row_q4_K_ref
    (same bytes as the reference's: tests/test_oracle_vs_ref.py, tests/golden)."""
    from oracle import port
    rng = np.random.default_rng(seed)
    w = rng.uniform(-1.0, 1.0, (M, K)).astype(np.float32)
    return port.quantize_q4_K(w)


def make_acts(N, K, seed=4321, adversarial=True):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((N, K)).astype(np.float32)
    if adversarial and K >= 512:
        x[0, :256] = 0.0                      # all-zero block
        x[-1, 256:512] = 0.0
        x[-1, 300] = 1e-30                    # single tiny non-zero
        x[0, 256] = -7.5                      # negative max
        x[0, 257] = 7.5                       # +/- tie: first occurrence wins
    return x


@pytest.fixture(scope="session")
def q4k():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import q4k_b200
    q4k_b200.load()        # raises if the CUDA library is missing: no fallback
    return q4k_b200


@pytest.fixture()
def exact_order(q4k):
    """decode path in the reference's AVX2 float order (bit-identical fp32); the default order is tolerance-only"""
    lib = q4k.load()
    lib.q4k_debug_set_gemv_order(1)
    yield lib
    lib.q4k_debug_set_gemv_order(0)
