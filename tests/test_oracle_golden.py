"""CPU: the oracle restatement against the committed vectors produced by the unmodified reference."""
import os

import numpy as np
import pytest

from oracle import port

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "q4k_ref_vectors.npz"))


def test_weight_quantizer_bytes():
    assert np.array_equal(port.quantize_q4_K(G["w"]), G["wq"])


def test_activation_quantizer_bytes():
    assert np.array_equal(port.quantize_q8_K(G["x"]), G["xq"])


def test_dequantize():
    assert np.array_equal(port.dequantize_q4_K(G["wq"]).view(np.uint32), G["w_deq"].view(np.uint32))


def test_mul_mat_avx2_order_bit_identical():
    got = port.mul_mat(G["wq"], G["x"], port.ORDER_AVX2)
    assert np.array_equal(got.view(np.uint32), G["y_avx2"].view(np.uint32))


def test_generic_and_neon_orders_close():
    f64 = port.mul_mat_f64(G["wq"], G["x"])
    scale = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
    for order in (port.ORDER_GENERIC, port.ORDER_NEON):
        got = port.mul_mat(G["wq"], G["x"], order)
        assert np.all(np.abs(got - G["y_generic"]) <= 1e-5 * np.maximum(np.abs(G["y_generic"]), scale))


def test_integer_stage_consistency():
    wq, xq = G["wq"], G["xq"]
    for n in range(xq.shape[0]):
        for m in range(wq.shape[0]):
            p, si, sm = port.vec_dot_partials(wq[m], xq[n])
            nb = p.shape[0]
            sc = np.zeros((nb, 8), np.int64); mn = np.zeros((nb, 8), np.int64)
            for b in range(nb):
                s, mm = port.unpack_scales_mins(wq[m][b * 144 + 4:b * 144 + 16])
                sc[b], mn[b] = s, mm
            assert np.array_equal((sc * p).sum(1), si)
            # independent numpy evaluation of p from the nibbles
            blk = wq[m].reshape(nb, 144)[:, 16:].reshape(nb, 4, 32)
            lv = np.stack([blk & 15, blk >> 4], axis=2).reshape(nb, 8, 32).astype(np.int64)
            q8 = xq[n].reshape(nb, 292)[:, 4:260].view(np.int8).reshape(nb, 8, 32).astype(np.int64)
            assert np.array_equal((lv * q8).sum(2), p)
            bs = xq[n].reshape(nb, 292)[:, 260:].copy().view(np.int16).reshape(nb, 8, 2).astype(np.int64).sum(2)
            assert np.array_equal((mn * bs).sum(1), sm)


def test_nearest_int_and_fp16():
    for f, want in [(0.5, 0), (1.5, 2), (2.5, 2), (-0.5, 0), (-1.5, -2), (126.5, 126), (127.49999, 127), (-127.5, -128)]:
        assert port.nearest_int(f) == want
    vals = np.concatenate([np.random.default_rng(0).standard_normal(5000).astype(np.float32) * s for s in (1e-8, 1e-4, 1, 1e3, 1e5)])
    ours = np.array([port.fp32_to_fp16_bits(v) for v in vals], np.uint16)
    assert np.array_equal(ours, vals.astype(np.float16).view(np.uint16))
    back = np.array([port.fp16_bits_to_fp32(h) for h in range(0, 0x7c00, 7)], np.float32)
    assert np.array_equal(back, np.arange(0, 0x7c00, 7, dtype=np.uint16).view(np.float16).astype(np.float32))


def test_empty_and_zero():
    assert port.quantize_q8_K(np.zeros((2, 256), np.float32)).sum() == 0
    assert port.mul_mat(G["wq"][:0], G["x"]).shape == (G["x"].shape[0], 0)
