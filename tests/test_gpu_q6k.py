"""GPU parity of the Q6_K x Q8_K decode path (csrc/gemv_q6k.cu through the C ABI: q6k_mul_mat, q6k_mul_mat_q8,
q6k_debug_partials) against the CPU oracle (oracle/q6k_oracle.c, pinned bit-exactly to the reference's
ggml_vec_dot_q6_K_q8_K) and the committed reference vectors.  Integer sums per superblock: bit-exact.  fp32 results:
within 1e-5 (tolerance of `assert_close`: |got - ref| <= 1e-5 * max(|ref|, rms of the column))."""
import os

import numpy as np
import pytest

from conftest import make_acts
from test_gpu_parity import assert_close, t

pytestmark = pytest.mark.gpu

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "q6k_ref_vectors.npz"))


def make_q6(M, K, seed=99, real=True):
    from oracle import port
    rng = np.random.default_rng(seed)
    if real:                      # quantize_row_q6_K_ref on U(-1,1), rows at three magnitudes
        w = (rng.uniform(-1, 1, (M, K)) * rng.choice([1e-2, 1.0, 30.0], (M, 1))).astype(np.float32)
        return port.quantize_q6_K(w)
    return port.random_q6_K(rng, M, K)


def test_golden_vectors(q4k):
    y = q4k.mul_mat_q6k(t(G["wq"]), t(G["x"])).cpu().numpy()
    from oracle import port
    assert_close(y, G["y_avx2"], port.mul_mat_q6_f64(G["wq"], G["x"]))
    yq = q4k.mul_mat_q6k_q8(t(G["wq"]), t(G["xq"])).cpu().numpy()
    assert np.array_equal(y.view(np.uint32), yq.view(np.uint32))           # fused quantizer == Q8_K input path, same bits


@pytest.mark.parametrize("M,K,N,real", [(33, 512, 1, True), (64, 2048, 2, True), (41, 5632, 3, False), (16, 4096, 8, False),
                                        (7, 256, 1, False), (9, 768, 4, True)])
def test_integer_sums_bit_exact(q4k, M, K, N, real):
    """the product kernel's per-superblock int32 sums == the reference's aux32 totals (oracle: ora_vec_dot_q6_partials)"""
    from oracle import port
    w = make_q6(M, K, real=real)
    xq = port.quantize_q8_K(make_acts(N, K))
    si, y = [a.cpu().numpy() for a in q4k.debug_partials_q6k(t(w), t(xq))]
    for n in range(N):
        for m in range(M):
            _, rsi = port.vec_dot_q6_partials(w[m], xq[n])
            assert np.array_equal(si[n, m], rsi), (n, m)
    y2 = q4k.mul_mat_q6k_q8(t(w), t(xq)).cpu().numpy()
    assert np.array_equal(y.view(np.uint32), y2.view(np.uint32))           # the dump build computes the same floats


@pytest.mark.parametrize("M,K,N", [(5632 // 8, 2048, 1), (256, 2048, 1), (2048, 5632, 1), (1000, 2048, 2), (130, 14336, 1),
                                   (257, 2048, 5), (64, 8192, 8), (3, 256, 1), (1, 256, 1), (5, 5632, 8)])
def test_mul_mat_within_tolerance(q4k, M, K, N):
    from oracle import port
    w = make_q6(M, K, seed=M + K, real=M * K <= 2048 * 2048)
    x = make_acts(N, K)
    y = q4k.mul_mat_q6k(t(w), t(x)).cpu().numpy()
    ref = port.mul_mat_q6(w, x, port.ORDER_AVX2)
    assert_close(y, ref, port.mul_mat_q6_f64(w, x) if M * N <= 4096 else ref.astype(np.float64))


def test_many_columns_and_long_k(q4k):
    """N > 8 runs eight columns per pass; K = 28672 with 8 columns does not fit shared memory and splits further"""
    import q4k_b200
    from oracle import port
    try:
        q4k_b200.debug_set("q6k_gemm_min", 0)              # keep this test on the decode kernel (N > 8 goes to the GEMM by default)
        for M, K, N in [(40, 1024, 19), (12, 28672, 8)]:
            w = make_q6(M, K, real=False)
            x = make_acts(N, K)
            y = q4k.mul_mat_q6k(t(w), t(x)).cpu().numpy()
            ref = port.mul_mat_q6(w, x, port.ORDER_AVX2)
            assert_close(y, ref, port.mul_mat_q6_f64(w, x))
    finally:
        q4k_b200.debug_set("q6k_gemm_min", 9)


def test_row_results_do_not_depend_on_position_or_batch(q4k):
    """a row gives the same bits alone, inside a larger matrix, at another row offset, and in any column of a batch"""
    w = make_q6(96, 2048, real=False)
    x = make_acts(4, 2048)
    full = q4k.mul_mat_q6k(t(w), t(x)).cpu().numpy()
    part = q4k.mul_mat_q6k(t(w[37:50]), t(x)).cpu().numpy()
    assert np.array_equal(full[:, 37:50].view(np.uint32), part.view(np.uint32))
    one = q4k.mul_mat_q6k(t(w), t(x[2:3])).cpu().numpy()
    assert np.array_equal(full[2:3].view(np.uint32), one.view(np.uint32))


def test_unaligned_rows_and_matrix_tail(q4k):
    """K = 5632: rows are 4620 B (12 mod 16), every row piece starts at another 16-byte phase; a matrix that ends off a
    16-byte boundary at the very end of its allocation must not be read past its last byte (compute-sanitizer clean by
    construction: the tail is copied with plain loads)."""
    import torch
    from oracle import port
    M, K = 37, 5632
    w = make_q6(M, K, real=False)
    x = make_acts(2, K)
    buf = torch.zeros(M * w.shape[1], dtype=torch.uint8, device="cuda")      # exact-size allocation
    buf.copy_(t(w).reshape(-1))
    y = q4k.mul_mat_q6k(buf.view(M, -1), t(x)).cpu().numpy()
    ref = port.mul_mat_q6(w, x, port.ORDER_AVX2)
    assert_close(y, ref, port.mul_mat_q6_f64(w, x))
    # a view that starts on an odd 2-byte phase inside a larger buffer
    big = torch.zeros(w.size + 64, dtype=torch.uint8, device="cuda")
    big[6:6 + w.size] = t(w).reshape(-1)
    y2 = q4k.mul_mat_q6k(big[6:6 + w.size].view(M, -1), t(x)).cpu().numpy()
    assert np.array_equal(y.view(np.uint32), y2.view(np.uint32))


def test_zero_activations_and_empty(q4k):
    import torch
    w = make_q6(8, 512, real=False)
    y = q4k.mul_mat_q6k(t(w), torch.zeros((2, 512), device="cuda")).cpu().numpy()
    assert np.array_equal(y, np.zeros((2, 8), np.float32))
    assert q4k.mul_mat_q6k(t(w[:0]), torch.zeros((1, 512), device="cuda")).shape == (1, 0)


def test_bad_arguments(q4k):
    import torch
    w = t(make_q6(4, 512, real=False))
    with pytest.raises(ValueError):
        q4k.mul_mat_q6k(w, torch.zeros((1, 256), device="cuda"))
    L = q4k.load()
    x = torch.zeros((1, 512), device="cuda")
    out = torch.zeros((1, 4), device="cuda")
    assert L.q6k_mul_mat(w.data_ptr() + 1, 4, 512, 420, x.data_ptr(), 1, 2048, out.data_ptr(), 16, 0, 0, 0) == -1   # odd pointer
    assert L.q6k_mul_mat(w.data_ptr(), 4, 500, 420, x.data_ptr(), 1, 2048, out.data_ptr(), 16, 0, 0, 0) == -1        # K % 256


def test_plan_mixes_q4k_and_q6k(q4k):
    """a graph plan holding Q4_K and Q6_K mul_mats (one Q4_K_M layer: q/k on Q4_K, v on Q6_K, all on one input) replays
    to the same bits as the direct calls"""
    import torch
    from conftest import make_weights
    K, N = 2048, 1
    wq, wk = t(make_weights(256, K, seed=1)), t(make_weights(64, K, seed=2))
    wv = t(make_q6(64, K, real=False))
    wd = t(make_q6(128, 5632, seed=5, real=False))
    x, x2 = t(make_acts(N, K)), t(make_acts(N, 5632))
    plan = q4k.GraphPlan()
    i0, i1 = plan.add(wq, N, 0), plan.add(wk, N, 0)
    i2 = plan.add(wv, N, 0, q6k=True)
    i3 = plan.add(wd, N, 1, q6k=True)
    plan.finalize()
    plan.input(0).copy_(x)
    plan.input(1).copy_(x2)
    for _ in range(2):
        plan.run()
    torch.cuda.synchronize()
    for idx, ref in ((i0, q4k.mul_mat(wq, x)), (i1, q4k.mul_mat(wk, x)), (i2, q4k.mul_mat_q6k(wv, x)), (i3, q4k.mul_mat_q6k(wd, x2))):
        assert torch.equal(plan.output(idx).view(torch.int32), ref.view(torch.int32))
