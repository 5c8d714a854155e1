"""GPU: the batched tensor-pipe decode kernel (gemv_q4k_wide.cu; 8 < N <= 64 by default, 32 columns per launch) against the oracle.
Integers from the DUMP instantiation of the product kernel must be exact; fp32 judged by tests/parity_util.py (1e-5)."""
import numpy as np
import pytest

from conftest import make_acts, make_weights
from parity_util import check

pytestmark = pytest.mark.gpu


def t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("M,K", [(64, 512), (300, 1024), (37, 2304), (16, 256), (1, 256), (129, 2560), (9600, 512), (40, 4096)])
@pytest.mark.parametrize("N", [2, 3, 5, 8, 9, 12, 13, 16, 17, 24, 32, 33, 40, 64])
def test_wide_within_tolerance_and_integers_exact(q4k, M, K, N):
    """ragged tiles (M % 16, odd tile counts), ragged rounds (K / 256 % 4), padded column tiles (N % 8), two column groups (N > 32)"""
    from oracle import port
    w, x = make_weights(M, K, seed=M + N), make_acts(N, K, seed=K + N)
    n0 = q4k.kernel_launches()
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    if N == 2 and K < 4096:       # two columns of a short K stay on the decode kernel (one launch, quantizer fused)
        assert q4k.kernel_launches() - n0 == 1
    elif N <= 32:
        assert q4k.kernel_launches() - n0 == 2                       # quantizer + batched decode kernel: not the GEMM, not decode passes
    if N <= 32:           # (above 32 columns q4k_mul_mat is the GEMM: tests/test_gpu_parity.py; this kernel's result is checked below)
        check(f"batched decode {M}x{K} N={N}", got, w, x)
    dst, p, si, sm = q4k.mul_mat_dump(t(w), t(x), q4k.DUMP_WIDE)         # N > 32: two column groups through the same kernel
    rp, rsi, rsm = port.partials_matrix(w, port.quantize_q8_K(x))
    assert np.array_equal(p.cpu().numpy(), rp) and np.array_equal(si.cpu().numpy(), rsi) and np.array_equal(sm.cpu().numpy(), rsm)
    if N <= 32 and not (N == 2 and K < 4096):
        assert np.array_equal(dst.cpu().numpy().view(np.uint32), got.view(np.uint32))     # the dump build computes the same floats
    else:
        check(f"batched decode (two column groups) {M}x{K} N={N}", dst.cpu().numpy(), w, x)


@pytest.mark.parametrize("N", [9, 32, 48])
def test_wide_integer_stage_is_exact(q4k, N):
    q4k.debug_set("midn_max", 64)
    try:
        _wide_integer_stage(q4k, N)
    finally:
        q4k.debug_set("midn_max", 32)


def _wide_integer_stage(q4k, N):
    """d = dmin = 1, activations that quantize to themselves, small scales: every float operation is exact, so the output is
    sumi - summ bit for bit whatever the accumulation order (same construction as tests/test_gpu_mma.py)."""
    from oracle import port
    rng = np.random.default_rng(70 + N)
    M, K = 100, 2560
    w = rng.integers(0, 256, (M, K // 256, 144), dtype=np.uint8)
    w[..., 0:2] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w[..., 2:4] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w[..., 4:16] &= 0x33
    w = w.reshape(M, -1)
    x = rng.integers(-20, 21, (N, K)).astype(np.float32)
    x[:, ::256] = -127.0
    _, rsi, rsm = port.partials_matrix(w, port.quantize_q8_K(x))
    want = rsi.sum(axis=2, dtype=np.int64) - rsm.sum(axis=2, dtype=np.int64)
    assert np.abs(want).max() < 2 ** 24
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    assert np.array_equal(got, want.astype(np.float32))


def test_wide_column_bits_do_not_depend_on_the_batch(q4k):
    """a column computed in a batch of 9, 32 or 64 (second column group) carries the same bits, and so does a row whether it is
    in the first or the last tile of a taller matrix"""
    K = 2048
    w, x = make_weights(200, K, seed=5), make_acts(64, K, seed=6)
    full = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    q4k.debug_set("midn_max", 64)                                  # 64 columns = two column groups of this kernel
    try:
        full = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    finally:
        q4k.debug_set("midn_max", 32)
    for n in (3, 8, 9, 24, 32):
        part = q4k.mul_mat(t(w), t(x[:n])).cpu().numpy()
        assert np.array_equal(part.view(np.uint32), full[:n].view(np.uint32))
    tail = q4k.mul_mat(t(w[150:]), t(x[40:56])).cpu().numpy()
    assert np.array_equal(tail.view(np.uint32), full[40:56, 150:].view(np.uint32))


def test_wide_strided_operands(q4k):
    """row strides wider than the rows on all three operands (what a ggml view hands over)"""
    import torch
    M, K, N = 80, 1024, 20
    w, x = make_weights(M, K, seed=11), make_acts(N, K, seed=12)
    nbw = K // 256 * 144
    wbuf = torch.zeros((M, nbw + 144), dtype=torch.uint8, device="cuda")
    wbuf[:, :nbw] = t(w)
    xbuf = torch.zeros((N, K + 64), dtype=torch.float32, device="cuda")
    xbuf[:, :K] = t(x)
    out = torch.full((N, M + 16), 7.0, dtype=torch.float32, device="cuda")
    lib = q4k.load()
    ws = torch.empty(q4k.workspace_bytes(K, N), dtype=torch.uint8, device="cuda")
    rc = lib.q4k_mul_mat(wbuf.data_ptr(), M, K, wbuf.stride(0), xbuf.data_ptr(), N, xbuf.stride(0) * 4, out.data_ptr(), out.stride(0) * 4,
                         ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    dense = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    o = out.cpu().numpy()
    assert np.array_equal(o[:, :M].view(np.uint32), dense.view(np.uint32)) and np.all(o[:, M:] == 7.0)
