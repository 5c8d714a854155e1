"""CPU, build container only: the oracle restatement against the reference itself (oracle/_ref), on
fresh random inputs and at the TinyLlama config.  Skipped where oracle/_ref is absent."""
import numpy as np
import pytest

from oracle import port, ref

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_quantizers_match_reference(seed):
    rng = np.random.default_rng(seed)
    w = (rng.uniform(-1, 1, (16, 1024)) * rng.choice([1e-3, 1, 50], (16, 1))).astype(np.float32)
    x = (rng.standard_normal((6, 1024)) * rng.choice([1e-4, 1, 1e4], (6, 1))).astype(np.float32)
    x[0, :256] = 0
    assert np.array_equal(port.quantize_q4_K(w), ref.quantize_q4_K(w))
    assert np.array_equal(port.quantize_q8_K(x), ref.quantize_q8_K(x))


def test_tinyllama_config_bit_identical():
    """configs[0]: Q4_K[5632x2048] x Q8_K vector (n=2048, nrc=1) on CPU ggml, whole op vs oracle (row sample for the port)."""
    rng = np.random.default_rng(1234)
    w = rng.uniform(-1, 1, (5632, 2048)).astype(np.float32)
    x = rng.standard_normal((1, 2048)).astype(np.float32)
    wq = ref.quantize_q4_K(w)
    y, _, _ = ref.mul_mat(wq, x, n_threads=4)
    got = port.mul_mat(wq, x, port.ORDER_AVX2)
    assert np.array_equal(got.view(np.uint32), y.view(np.uint32))
    assert np.array_equal(port.quantize_q4_K(w[:64]), wq[:64])


def test_reference_takes_avx2_branch():
    assert ref.has_avx2()
