"""CPU: the Q6_K x Q8_K oracle restatement (oracle/q6k_oracle.c) against vectors produced by the unmodified reference
(tests/golden/q6k_ref_vectors.npz, made by tests/golden/make_golden.py) and, where oracle/_ref is present, against the
reference itself on fresh inputs."""
import os

import numpy as np
import pytest

from oracle import port, ref

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "q6k_ref_vectors.npz"))


def test_weight_quantizer_bytes():
    assert np.array_equal(port.quantize_q6_K(G["w"]), G["wq"])


def test_dequantize():
    assert np.array_equal(port.dequantize_q6_K(G["wq"]).view(np.uint32), G["w_deq"].view(np.uint32))


def test_mul_mat_avx2_order_bit_identical():
    got = port.mul_mat_q6(G["wq"], G["x"], port.ORDER_AVX2)
    assert np.array_equal(got.view(np.uint32), G["y_avx2"].view(np.uint32))


def test_generic_order_bit_identical():
    got = port.mul_mat_q6(G["wq"], G["x"], port.ORDER_GENERIC)
    assert np.array_equal(got.view(np.uint32), G["y_generic"].view(np.uint32))


def test_integer_stage_consistency():
    """p / sumi against an independent numpy evaluation of the 6-bit levels (ggml-quants.c:1772-1779)."""
    wq, xq = G["wq"], G["xq"]
    nb = wq.shape[1] // 210
    blk = wq.reshape(-1, nb, 210)
    ql = blk[:, :, :128].reshape(-1, nb, 2, 2, 32).astype(np.int64)       # [half][which 32][l]
    qh = blk[:, :, 128:192].reshape(-1, nb, 2, 32).astype(np.int64)
    sc = blk[:, :, 192:208].view(np.int8).astype(np.int64)
    lv = np.zeros((wq.shape[0], nb, 2, 4, 32), np.int64)
    for t in range(4):
        nib = (ql[:, :, :, t & 1] >> 4) if t & 2 else (ql[:, :, :, t & 1] & 15)
        lv[:, :, :, t] = (nib | (((qh >> (2 * t)) & 3) << 4)) - 32
    lv = lv.reshape(wq.shape[0], nb, 16, 16)
    for n in range(xq.shape[0]):
        q8 = xq[n].reshape(nb, 292)[:, 4:260].view(np.int8).reshape(nb, 16, 16).astype(np.int64)
        for m in range(wq.shape[0]):
            p, si = port.vec_dot_q6_partials(wq[m], xq[n])
            assert np.array_equal((lv[m] * q8).sum(2), p)
            assert np.array_equal((sc[m] * p).sum(1), si)
            # the bsums identity the CUDA kernel uses: sum (q6 - 32) q8 = sum q6 q8 - 32 * bsums
            bs = xq[n].reshape(nb, 292)[:, 260:].copy().view(np.int16).reshape(nb, 16).astype(np.int64)
            assert np.array_equal(((lv[m] + 32) * q8).sum(2) - 32 * bs, p)


def test_random_valid_blocks_and_empty():
    rng = np.random.default_rng(3)
    w = port.random_q6_K(rng, 6, 512)
    x = rng.standard_normal((2, 512)).astype(np.float32)
    y = port.mul_mat_q6(w, x)
    f64 = port.mul_mat_q6_f64(w, x)
    assert np.all(np.isfinite(y)) and np.allclose(y, f64, rtol=1e-4, atol=1e-4 * np.abs(f64).max())
    assert port.mul_mat_q6(w[:0], x).shape == (2, 0)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed", [0, 1])
def test_against_reference_fresh_inputs(seed):
    rng = np.random.default_rng(seed)
    w = (rng.uniform(-1, 1, (10, 1024)) * rng.choice([1e-3, 1, 50], (10, 1))).astype(np.float32)
    w[2, 256:512] = 0
    x = (rng.standard_normal((3, 1024)) * rng.choice([1e-4, 1, 1e4], (3, 1))).astype(np.float32)
    wq = ref.quantize_q6_K(w)
    assert np.array_equal(port.quantize_q6_K(w), wq)
    xq = ref.quantize_q8_K(x)
    for W in (wq, port.random_q6_K(rng, 6, 1024)):
        for m in range(W.shape[0]):
            for n in range(3):
                assert port.vec_dot_q6(W[m], xq[n], port.ORDER_AVX2).view(np.uint32) == ref.vec_dot_q6(W[m], xq[n]).view(np.uint32)
                assert port.vec_dot_q6(W[m], xq[n], port.ORDER_GENERIC).view(np.uint32) == \
                    ref.vec_dot_q6(W[m], xq[n], generic=True).view(np.uint32)
