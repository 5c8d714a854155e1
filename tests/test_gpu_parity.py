"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bit-exact for Q8_K bytes and the int32 partial sums (dumped by the PRODUCT kernels); fp32 outputs bit-identical to the
AVX2-order oracle where the kernel runs the exact order, else judged by tests/parity_util.py (literal 1e-5 relative statistic,
no worse than the reference's own variant spread, scaled 1e-5 bound)."""
import numpy as np
import pytest

from conftest import make_acts, make_weights
from parity_util import check, log, rel_err, TOL

pytestmark = pytest.mark.gpu


def t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def assert_close(got, ref, f64, tol=TOL, name="q6k"):
    """The scaled bound alone, |got - ref| <= tol * max(|ref|, rms of the column), with the literal statistic printed — for the
    comparisons that have no generic-order companion at hand (Q6_K, plans).  Q4_K comparisons use parity_util.check."""
    r = rel_err(got, ref)
    log(name, {"n": int(got.size), "max_rel_vs_avx2": float(r.max(initial=0.0)), "n_above_1e-5_vs_avx2": int((r > tol).sum())})
    rms = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
    bound = tol * np.maximum(np.abs(ref), rms)
    bad = np.abs(got.astype(np.float64) - ref.astype(np.float64)) > bound
    assert not bad.any(), f"{bad.sum()} elements off; worst {np.abs(got - ref).max()}"


@pytest.mark.parametrize("N,K", [(1, 256), (3, 2048), (8, 4096), (512, 512)])
def test_quantize_q8k_bit_exact(q4k, N, K):
    from oracle import port
    x = make_acts(N, K)
    x[0, -1] = 0.5 / 127 * np.abs(x[0, -256:]).max()      # lands near a .5 tie after scaling
    got = q4k.quantize_q8_K(t(x)).cpu().numpy()
    assert np.array_equal(got, port.quantize_q8_K(x))


def test_quantize_q8k_denormals_and_ties(q4k):
    from oracle import port
    x = np.zeros((2, 512), np.float32)
    x[0, :256] = np.float32(1e-40)                          # denormals
    x[0, 256:] = np.tile(np.array([1.0, -1.0], np.float32), 128)
    x[1, :256] = np.linspace(-127, 127, 256).astype(np.float32) + 0.5
    x[1, 256:] = np.float32(3.4e38)
    got = q4k.quantize_q8_K(t(x)).cpu().numpy()
    assert np.array_equal(got, port.quantize_q8_K(x))


DECODE_KERNELS = ["DUMP_DISPATCH", "DUMP_TENSOR_PIPE", "DUMP_BULK", "DUMP_BULK_EXACT"]


@pytest.mark.parametrize("kernel", DECODE_KERNELS)
@pytest.mark.parametrize("M,K,N", [(33, 512, 1), (64, 2048, 2), (40, 1024, 5), (16, 4096, 8), (100, 2048, 1), (50, 14336, 1), (21, 8448, 3), (37, 28672, 1)])
def test_int32_partials_bit_exact_from_the_product_decode_kernels(q4k, kernel, M, K, N):
    """Per-sub-block p[j], sumi and summ of EVERY (column, row, superblock), written by the DUMP instantiation of the product
    kernel itself (same code, the MMA C fragments / dp4a sums it really computed), against the generic reference's integers
    (L/ggml/src/ggml-cpu/quants.c:575-611); the fp32 results of that launch equal the plain launch's bit for bit."""
    from oracle import port
    w, x = make_weights(M, K), make_acts(N, K)
    rp, rsi, rsm = port.partials_matrix(w, port.quantize_q8_K(x))
    dst, p, si, sm = q4k.mul_mat_dump(t(w), t(x), getattr(q4k, kernel))
    assert np.array_equal(p.cpu().numpy(), rp)
    assert np.array_equal(si.cpu().numpy(), rsi)
    assert np.array_equal(sm.cpu().numpy(), rsm)
    if kernel == "DUMP_DISPATCH":
        assert np.array_equal(dst.cpu().numpy().view(np.uint32), q4k.mul_mat(t(w), t(x)).cpu().numpy().view(np.uint32))


@pytest.mark.parametrize("M,K", [(33, 512), (70, 8192), (20, 28672)])
def test_int32_partials_bit_exact_from_the_large_matrix_kernel(q4k, M, K):
    from oracle import port
    w, x = make_weights(M, K), make_acts(1, K)
    rp, rsi, rsm = port.partials_matrix(w, port.quantize_q8_K(x))
    dst, p, si, sm = q4k.mul_mat_dump(t(w), t(x), q4k.DUMP_BIG)
    assert np.array_equal(p.cpu().numpy(), rp) and np.array_equal(si.cpu().numpy(), rsi) and np.array_equal(sm.cpu().numpy(), rsm)


@pytest.fixture()
def gemm_path(q4k):
    """8 < N <= 64 through the tcgen05 GEMM (as for N > 64) instead of the batched decode kernel (tests/test_gpu_wide.py)"""
    q4k.debug_set("midn_max", 0)
    yield
    q4k.debug_set("midn_max", 32)


@pytest.mark.parametrize("M,K,N", [(128, 512, 16), (200, 2048, 40), (130, 1024, 130), (64, 14336, 32), (300, 4096, 9)])
def test_int32_sums_bit_exact_from_the_gemm_epilogue(q4k, gemm_path, M, K, N):
    """The prefill GEMM never forms p[j] (the scales ride in the A operand), but its epilogue reads sumi = 8 hi + lo and summ
    of every (row, column, superblock) out of TMEM: the DUMP instantiation writes exactly those integers."""
    from oracle import port
    w, x = make_weights(M, K), make_acts(N, K)
    _, rsi, rsm = port.partials_matrix(w, port.quantize_q8_K(x))
    dst, p, si, sm = q4k.mul_mat_dump(t(w), t(x), q4k.DUMP_GEMM)
    assert p is None
    assert np.array_equal(si.cpu().numpy(), rsi) and np.array_equal(sm.cpu().numpy(), rsm)
    if N > 12:                                              # q4k_mul_mat serves N <= 12 with the decode kernel: other float order
        assert np.array_equal(dst.cpu().numpy().view(np.uint32), q4k.mul_mat(t(w), t(x)).cpu().numpy().view(np.uint32))
    check(f"gemm (dump instantiation) {M}x{K} N={N}", dst.cpu().numpy(), w, x)


@pytest.mark.parametrize("M,K,N", [(5632, 2048, 1), (100, 512, 1), (257, 2048, 2), (64, 14336, 1), (1, 256, 1)])
def test_gemv_bit_identical_to_avx2_reference_order(q4k, exact_order, M, K, N):
    from oracle import port
    w, x = make_weights(M, K), make_acts(N, K)
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    ref = port.mul_mat(w, x, port.ORDER_AVX2)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("M,K,N", [(300, 2048, 3), (129, 1024, 4), (77, 2048, 5), (64, 4096, 8), (50, 512, 7), (40, 512, 6)])
def test_gemv_batch_within_tolerance(q4k, M, K, N):
    from oracle import port
    w, x = make_weights(M, K), make_acts(N, K)
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    check(f"gemv {M}x{K} N={N}", got, w, x)


@pytest.mark.parametrize("N", [9, 32, 100])
def test_mul_mat_large_batch(q4k, N):
    from oracle import port
    M, K = 128, 1024
    w, x = make_weights(M, K), make_acts(N, K)
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    check(f"mul_mat {M}x{K} N={N}", got, w, x)


@pytest.mark.parametrize("M,K,N", [(5632, 2048, 1), (2048, 5632, 1), (259, 2048, 2), (64, 8192, 4), (33, 1024, 8)])
def test_fused_quantize_equals_two_kernel_path(q4k, exact_order, M, K, N):
    """q4k_mul_mat fuses quantize_row_q8_K into the GEMV launch for decode shapes; the result must be bit-identical
    to quantize kernel + GEMV (and, for N <= 2, to the AVX2-order oracle), with and without PDL launches."""
    from oracle import port
    lib = q4k.load()
    w, x = make_weights(M, K), make_acts(N, K)
    wd, xd = t(w), t(x)
    try:
        lib.q4k_debug_set_fuse_limit(0)
        two = q4k.mul_mat(wd, xd).cpu().numpy()
        lib.q4k_debug_set_fuse_limit(1 << 20)
        fused = q4k.mul_mat(wd, xd).cpu().numpy()
        lib.q4k_debug_set_pdl(0)
        nopdl = q4k.mul_mat(wd, xd).cpu().numpy()
    finally:
        lib.q4k_debug_set_fuse_limit(16384)
        lib.q4k_debug_set_pdl(1)
    assert np.array_equal(two.view(np.uint32), fused.view(np.uint32))
    assert np.array_equal(two.view(np.uint32), nopdl.view(np.uint32))
    ref = port.mul_mat(w, x, port.ORDER_AVX2)
    if N <= 2:
        assert np.array_equal(fused.view(np.uint32), ref.view(np.uint32))
    else:
        check(f"fused {M}x{K} N={N}", fused, w, x)


def test_back_to_back_pdl_launches_see_fresh_activations(q4k, exact_order):
    """Chained mul_mats on one stream (each reads what the previous wrote) under programmatic dependent launch."""
    import torch
    from oracle import port
    K = 2048
    w1, w2 = make_weights(K, K, seed=5), make_weights(K, K, seed=6)
    x = make_acts(1, K, adversarial=False)
    w1d, w2d, xd = t(w1), t(w2), t(x)
    for _ in range(3):
        y1 = q4k.mul_mat(w1d, xd)
        y2 = q4k.mul_mat(w2d, y1)
        y3 = q4k.mul_mat(w1d, y2)
    r1 = port.mul_mat(w1, x, port.ORDER_AVX2)
    r2 = port.mul_mat(w2, r1, port.ORDER_AVX2)
    r3 = port.mul_mat(w1, r2, port.ORDER_AVX2)
    assert np.array_equal(y3.cpu().numpy().view(np.uint32), r3.view(np.uint32))


@pytest.mark.parametrize("M,K,N", [(128, 256, 16), (256, 2048, 128), (130, 512, 9), (300, 1024, 200), (64, 4096, 130), (512, 5632, 64)])
def test_gemm_tensor_core_within_tolerance(q4k, gemm_path, M, K, N):
    """N > 8 takes the tcgen05 kind::i8 GEMM (exact int32 sumi / summ per superblock in TMEM, fp32 scales in the
    epilogue); fp32 results within 1e-5 of the AVX2-order oracle, ragged M / N tiles included."""
    from oracle import port
    w, x = make_weights(M, K), make_acts(N, K)
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    check(f"gemm {M}x{K} N={N}", got, w, x)


def test_gemm_integer_stage_is_exact(q4k, gemm_path):
    """With d = dmin = 1 in every weight block and activations that quantize to themselves (d_x = 1), the fp32 output
    is exactly sumi - summ: any error in the tensor-core integer stage (scale split, min accumulator) shows."""
    from oracle import port
    rng = np.random.default_rng(5)
    M, K, N = 128, 512, 32
    w = rng.integers(0, 256, (M, K // 256, 144), dtype=np.uint8)
    w[..., 0:2] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w[..., 2:4] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w = w.reshape(M, -1)
    x = rng.integers(-127, 128, (N, K)).astype(np.float32)
    x[:, ::256] = -127.0                                    # block max -> iscale = 1, d = 1
    xq = port.quantize_q8_K(x)
    want = np.zeros((N, M), np.int64)
    for n in range(N):
        for m in range(M):
            _, si, sm = port.vec_dot_partials(w[m], xq[n])
            want[n, m] = int(si.sum()) - int(sm.sum())
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    assert np.array_equal(got.astype(np.int64), want) and np.array_equal(got, want.astype(np.float32))


@pytest.mark.parametrize("ksplit", [1, 2, 3, 8])
def test_gemm_split_k_integer_stage_is_exact_and_deterministic(q4k, gemm_path, ksplit):
    """The persistent grid forced to tiles x ksplit CTAs: tiles are split between CTAs (later parts parked in the workspace, the
    part with the first superblocks adds them in K order; tests/test_gpu_gemm_streamk.py has the launches with more tiles than SMs).
    Same exact-float construction as above at K = 2048: any lost / doubled superblock or a race on the flags shows as a
    wrong integer; two runs give identical bits."""
    from oracle import port
    q4k.debug_set("gemm_ksplit", ksplit)
    rng = np.random.default_rng(11)
    M, K, N = 200, 2048, 40
    w = rng.integers(0, 256, (M, K // 256, 144), dtype=np.uint8)
    w[..., 0:2] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w[..., 2:4] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w = w.reshape(M, -1)
    x = rng.integers(-127, 128, (N, K)).astype(np.float32)
    x[:, ::256] = -127.0
    xq = port.quantize_q8_K(x)
    want = np.zeros((N, M), np.int64)
    for n in range(N):
        for m in range(M):
            _, si, sm = port.vec_dot_partials(w[m], xq[n])
            want[n, m] = int(si.sum()) - int(sm.sum())
    assert np.abs(want).max() < 2 ** 24
    try:
        got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
        again = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    finally:
        q4k.debug_set("gemm_ksplit", 0)
    assert np.array_equal(got, want.astype(np.float32))
    assert np.array_equal(got.view(np.uint32), again.view(np.uint32))


@pytest.mark.parametrize("N,K", [(16, 512), (130, 2048), (40, 4096)])
def test_prefill_reuses_the_quantized_activations(q4k, gemm_path, N, K):
    """q/k/v of a layer are three mul_mats on one src1: q4k_mul_mat_reuse runs the GEMM on the quantized copy the first call
    left in the workspace (one launch instead of two) and gives the bits of a full q4k_mul_mat; N <= 12 is declined."""
    import torch
    lib = q4k.load()
    w1, w2 = t(make_weights(256, K, seed=1)), t(make_weights(96, K, seed=2))
    x = t(make_acts(N, K))
    ws = torch.empty(q4k.workspace_bytes(K, N), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    want = q4k.mul_mat(w2, x, workspace=ws.clone())
    q4k.mul_mat(w1, x, workspace=ws)                        # leaves the quantized copy of x in ws
    got = torch.zeros((N, 96), device="cuda")
    n0 = q4k.kernel_launches()
    assert lib.q4k_mul_mat_reuse(w2.data_ptr(), 96, K, w2.stride(0), N, got.data_ptr(), 96 * 4, ws.data_ptr(), ws.numel(), st) == 0
    assert q4k.kernel_launches() == n0 + 1
    torch.cuda.synchronize()
    assert torch.equal(got.view(torch.int32), want.view(torch.int32))
    assert lib.q4k_mul_mat_reuse(w2.data_ptr(), 96, K, w2.stride(0), 8, got.data_ptr(), 96 * 4, ws.data_ptr(), ws.numel(), st) == -5


def test_vec_dot_signature_bit_identical(q4k):
    import torch
    from oracle import port
    K = 2048
    w = make_weights(4, K)
    xq = port.quantize_q8_K(make_acts(1, K))
    s = torch.zeros(1, device="cuda")
    for m in range(4):
        q4k.vec_dot_q4_K_q8_K(K, s, 0, t(w[m]), 0, t(xq[0]), 0, 1)
        assert np.float32(s.item()).view(np.uint32) == port.vec_dot(w[m], xq[0], port.ORDER_AVX2).view(np.uint32)
    with pytest.raises(q4k.Q4KError):
        q4k.vec_dot_q4_K_q8_K(K, s, 0, t(w[0]), 0, t(xq[0]), 0, 2)      # nrc != 1: the reference asserts


def test_error_codes_and_empty(q4k):
    import torch
    w = t(make_weights(4, 256))
    with pytest.raises((q4k.Q4KError, ValueError)):
        q4k.mul_mat(w, torch.zeros((1, 300), device="cuda"))
    out = q4k.mul_mat(t(make_weights(4, 256)[:0]), torch.zeros((1, 256), device="cuda"))
    assert out.shape == (1, 0)
    lib = q4k.load()
    assert lib.q4k_mul_mat(w.data_ptr(), 4, 256, 144, 0, 1, 1024, 0, 16, 0, 0, 0) in (-2, -3)
    assert lib.q4k_quantize_q8_K(0, 0, 100, 1, 400, 0) == -1


def test_launch_counter_moves(q4k):
    import torch
    n0 = q4k.kernel_launches()
    q4k.mul_mat(t(make_weights(8, 256)), torch.ones((1, 256), device="cuda"))
    assert q4k.kernel_launches() >= n0 + 1
