"""GPU parity of the Q6_K prefill GEMM (csrc/gemm_q6k_tc.cu: tcgen05 kind::f16, two exact fp16 pieces per weight) through the
C ABI (q6k_mul_mat at N > 8, q6k_mul_mat_dump) against the CPU oracle (oracle/q6k_oracle.c, pinned bit-exactly to the
reference's ggml_vec_dot_q6_K_q8_K, L/ggml/src/ggml-cpu/quants.c:795-848).
Integer sums per superblock, read out of tensor memory by the product kernel's dump instantiation: bit-exact — this is the
test that the fp32 accumulator of the fp16 tensor-core path never rounds.  fp32 results: within 1e-5
(|got - ref| <= 1e-5 * max(|ref|, rms of the column), `assert_close`)."""
import numpy as np
import pytest

from conftest import make_acts
from test_gpu_parity import assert_close, t
from test_gpu_q6k import make_q6

pytestmark = pytest.mark.gpu


def q6_scaled_ints(w):
    """block_q6_K bytes [M, nb*210] -> int32 [M, nb, 256] = sc[y / 16] * (q6[y] - 32) in element order
    (the layout dequantize_row_q6_K walks, L/ggml/src/ggml-quants.c:1762-1795)."""
    M = w.shape[0]
    b = w.reshape(M, -1, 210)
    ql, qh, sc = b[:, :, :128].astype(np.int32), b[:, :, 128:192].astype(np.int32), b[:, :, 192:208].view(np.int8).astype(np.int32)
    out = np.zeros(b.shape[:2] + (256,), np.int32)
    for n in range(2):
        l, h = ql[:, :, 64 * n:64 * n + 64], qh[:, :, 32 * n:32 * n + 32]
        q = [(l[:, :, :32] & 15) | ((h & 3) << 4), (l[:, :, 32:] & 15) | (((h >> 2) & 3) << 4),
             (l[:, :, :32] >> 4) | (((h >> 4) & 3) << 4), (l[:, :, 32:] >> 4) | (((h >> 6) & 3) << 4)]
        for e in range(4):
            out[:, :, 128 * n + 32 * e:128 * n + 32 * e + 32] = q[e] - 32
    return out * np.repeat(sc, 16, axis=2)


def sumi_ref(w, xq):
    """[N, M, nb] int64: sum over each superblock of sc * (q6 - 32) * q8"""
    a = q6_scaled_ints(w).astype(np.int64)
    q8 = xq.reshape(xq.shape[0], -1, 292)[:, :, 4:260].view(np.int8).astype(np.int64)
    return np.einsum("mbk,nbk->nmb", a, q8)


@pytest.mark.parametrize("M,K,N,real", [(128, 512, 64, False), (200, 2048, 130, False), (77, 5632, 33, False), (16, 256, 1, False),
                                        (129, 768, 65, True), (1, 256, 128, False), (300, 1280, 257, True)])
def test_integer_sums_from_tensor_memory_are_exact(q4k, M, K, N, real):
    from oracle import port
    w = make_q6(M, K, seed=M + N, real=real)
    x = make_acts(N, K)
    xq = port.quantize_q8_K(x)
    si, y = [a.cpu().numpy() for a in q4k.mul_mat_q6k_dump(t(w), t(x))]
    assert np.array_equal(si.astype(np.int64), sumi_ref(w, xq))
    for n, m in [(0, 0), (N - 1, M - 1), (N // 2, M // 3)]:                 # the numpy restatement above against the oracle proper
        assert np.array_equal(si[n, m], port.vec_dot_q6_partials(w[m], xq[n])[1])
    ref = port.mul_mat_q6(w, x, port.ORDER_AVX2)
    assert_close(y, ref, port.mul_mat_q6_f64(w, x) if M * N <= 4096 else ref.astype(np.float64), name="q6k_gemm_dump")


def test_saturated_blocks_stay_below_2_pow_24(q4k):
    """every weight at q6 = 0 (a = -32) or 63 (a = 31) with scales 127 / -128 / 15 / -16 and activations at +-127: the largest
    sums a superblock can produce (|S_lo| = 256 * 480 * 127 = 15 605 760 < 2^24) come back exact"""
    from oracle import port
    M, K, N = 128, 1024, 64
    nb = K // 256
    w = np.zeros((M, nb, 210), np.uint8)
    scales = np.array([127, -128, 15, -16], np.int8).view(np.uint8)
    for m in range(M):
        for b in range(nb):
            if (m + b) & 1:
                w[m, b, :192] = 0xFF                                     # q6 = 63
            w[m, b, 192:208] = scales[(m // 2 + b) % 4]
    w[:, :, 208:210] = np.array([1e-3], np.float16).view(np.uint8)
    w = w.reshape(M, nb * 210)
    x = np.ones((N, K), np.float32)
    x[1::2] *= -1.0
    x[2] = np.where(np.arange(K) % 2 == 0, 1.0, -1.0)
    xq = port.quantize_q8_K(x)
    si, _ = [a.cpu().numpy() for a in q4k.mul_mat_q6k_dump(t(w), t(x))]
    ref = sumi_ref(w, xq)
    assert np.abs(ref).max() == 256 * 32 * 128 * 127
    assert np.array_equal(si.astype(np.int64), ref)


@pytest.mark.parametrize("M,K,N", [(256, 2048, 512), (2048, 5632, 64), (1000, 2048, 33), (130, 14336, 96), (64, 8192, 40)])
def test_default_dispatch_takes_the_gemm_and_is_within_tolerance(q4k, M, K, N):
    import q4k_b200
    from oracle import port
    w = make_q6(M, K, seed=M + K, real=M * K <= 2048 * 2048)
    x = make_acts(N, K)
    rows = np.arange(M) if M * N <= 1 << 16 else np.unique(np.r_[0:64, M // 2:M // 2 + 64, M - 64:M])
    before = q4k_b200.kernel_launches()
    y = q4k.mul_mat_q6k(t(w), t(x)).cpu().numpy()
    assert q4k_b200.kernel_launches() - before == 2                            # quantizer + GEMM
    ref = port.mul_mat_q6(w[rows], x, port.ORDER_AVX2)
    assert_close(y[:, rows], ref, ref.astype(np.float64), name="q6k_gemm")
    # the same bits as the dump instantiation, and a row's bits do not depend on where it sits in the matrix
    _, yd = q4k.mul_mat_q6k_dump(t(w), t(x))
    assert np.array_equal(y.view(np.uint32), yd.cpu().numpy().view(np.uint32))
    y2 = q4k.mul_mat_q6k(t(np.ascontiguousarray(w[5:])), t(x)).cpu().numpy()
    assert np.array_equal(y[:, 5:].view(np.uint32), y2.view(np.uint32))


def test_split_k_keeps_the_integers_and_the_tolerance(q4k):
    import q4k_b200
    from oracle import port
    M, K, N = 256, 14336, 64                           # 2 row tiles: the launcher splits K over 2..4 CTAs per tile
    w = make_q6(M, K, real=False)
    x = make_acts(N, K)
    xq = port.quantize_q8_K(x)
    ref = port.mul_mat_q6(w, x, port.ORDER_AVX2)
    try:
        for ks in (1, 2, 3):
            q4k_b200.debug_set("gemm_ksplit", ks)
            si, y = [a.cpu().numpy() for a in q4k.mul_mat_q6k_dump(t(w), t(x))]
            assert np.array_equal(si.astype(np.int64), sumi_ref(w, xq)), ks
            assert_close(y, ref, ref.astype(np.float64), name=f"q6k_gemm_ksplit{ks}")
    finally:
        q4k_b200.debug_set("gemm_ksplit", 0)


def test_unaligned_weight_pointer_and_padded_row_stride(q4k):
    """210-byte blocks: odd block counts put every other row at 2 (mod 4); a view into a larger buffer moves the base by 2"""
    import torch
    from oracle import port
    M, K, N = 70, 768, 48
    w = make_q6(M, K, real=False)
    x = make_acts(N, K)
    ref = port.mul_mat_q6(w, x, port.ORDER_AVX2)
    stride = w.shape[1] + 6
    buf = torch.zeros(2 + M * stride + 64, dtype=torch.uint8, device="cuda")
    view = buf[2:2 + M * stride].view(M, stride)[:, :w.shape[1]]
    view.copy_(t(w))
    assert view.data_ptr() % 4 == 2
    L = q4k.load()
    xt = t(x)
    out = torch.empty((N, M), dtype=torch.float32, device="cuda")
    nbytes = L.q6k_mul_mat_workspace_bytes(K, N)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    assert L.q6k_mul_mat(view.data_ptr(), M, K, stride, xt.data_ptr(), N, K * 4, out.data_ptr(), M * 4, ws.data_ptr(), nbytes,
                         torch.cuda.current_stream().cuda_stream) == 0
    y = out.cpu().numpy()
    assert_close(y, ref, port.mul_mat_q6_f64(w, x), name="q6k_gemm_unaligned")
    y0 = q4k.mul_mat_q6k(t(w), xt).cpu().numpy()
    assert np.array_equal(y.view(np.uint32), y0.view(np.uint32))


def test_gemm_threshold_knob_and_workspace(q4k):
    import q4k_b200
    import torch
    L = q4k_b200.load()
    M, K, N = 64, 512, 40
    w, x = t(make_q6(M, K, real=False)), t(make_acts(N, K))
    y_gemm = q4k.mul_mat_q6k(w, x)
    try:
        q4k_b200.debug_set("q6k_gemm_min", 0)                                  # never: eight columns per decode pass
        assert L.q6k_mul_mat_workspace_bytes(K, N) == N * (K // 256) * 292
        y_gemv = q4k.mul_mat_q6k(w, x)
    finally:
        q4k_b200.debug_set("q6k_gemm_min", 9)
    assert torch.allclose(y_gemm, y_gemv, rtol=1e-4, atol=1e-4)
    assert L.q6k_mul_mat_workspace_bytes(K, N) >= L.q6k_gemm_workspace_bytes(K, N) > 128 * K * 2
    out = torch.empty((N, M), dtype=torch.float32, device="cuda")
    ws = torch.empty(1024, dtype=torch.uint8, device="cuda")
    assert L.q6k_mul_mat(w.data_ptr(), M, K, w.stride(0), x.data_ptr(), N, K * 4, out.data_ptr(), M * 4, ws.data_ptr(), 1024, 0) == -3
