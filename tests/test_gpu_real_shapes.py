"""GPU parity at BASELINE.json's REAL shapes, through the default dispatch (no forced knobs): C3 (5632x2048x512), C4 (4096x14336 and
14336x4096 at N = 1 / 8 / 32 / 512: decode kernel with cluster split-K, the batched decode kernel, the GEMM with its persistent
grid) and C5 (28672x8192 and 8192x28672 at N = 1 / 8: the large-matrix kernel, the separate quantizer).

Every row of the matrix is streamed by the kernel under test; a fixed sample of 256 rows — the first tile, a tile in the middle,
the last rows of the matrix and 64 rows scattered over it — holds weights really quantized by quantize_row_q4_K (the oracle's
restatement, byte-identical to the reference's) and is compared with the oracle: fp32 by tests/parity_util.py, and for N <= 8 also
the per-sub-block integers dumped by the product kernel on that sample.  The other rows hold random valid block_q4_K bytes
(any byte pattern with finite fp16 d / dmin is a valid block)."""
import numpy as np
import pytest

from parity_util import check

pytestmark = pytest.mark.gpu


def t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def sample_rows(M, rng):
    mid = (M // 2) // 128 * 128
    rows = set(range(0, 64)) | set(range(mid, mid + 64)) | set(range(M - 64, M))
    rows |= set(int(r) for r in rng.choice(M, 64, replace=False))
    return np.array(sorted(rows))


def build_case(M, K, N, seed):
    from oracle import port
    rng = np.random.default_rng(seed)
    nb = K // 256
    w = rng.integers(0, 256, (M, nb, 144), dtype=np.uint8)
    w[..., 1] = rng.integers(0x1c, 0x2c, (M, nb), dtype=np.uint8)       # d    in [2^-8, 2^-4): finite, sane magnitudes
    w[..., 3] = rng.integers(0x1c, 0x2c, (M, nb), dtype=np.uint8)       # dmin likewise
    w = w.reshape(M, nb * 144)
    rows = sample_rows(M, rng)
    w[rows] = port.quantize_q4_K(rng.uniform(-1.0, 1.0, (len(rows), K)).astype(np.float32))
    x = rng.standard_normal((N, K)).astype(np.float32)
    x[0, :256] = 0.0                                                    # an all-zero activation block
    x[N - 1, K - 256] = -9.0                                            # a negative block maximum
    return w, x, rows


REAL = [(5632, 2048, 512),
        (4096, 14336, 1), (4096, 14336, 8), (4096, 14336, 32), (4096, 14336, 512),
        (14336, 4096, 1), (14336, 4096, 8), (14336, 4096, 32), (14336, 4096, 512),
        (28672, 8192, 1), (28672, 8192, 8), (8192, 28672, 1), (8192, 28672, 8)]


@pytest.mark.parametrize("M,K,N", REAL)
def test_real_shape_default_dispatch(q4k, M, K, N):
    import torch
    w, x, rows = build_case(M, K, N, seed=M + K + N)
    wd, xd = t(w), t(x)
    n0 = q4k.kernel_launches()
    got = q4k.mul_mat(wd, xd)
    torch.cuda.synchronize()
    assert q4k.kernel_launches() > n0
    again = q4k.mul_mat(wd, xd).cpu().numpy()
    got = got.cpu().numpy()
    assert np.array_equal(got.view(np.uint32), again.view(np.uint32)), "two runs of the same launch differ"
    assert np.isfinite(got).all()
    check(f"real shape {M}x{K} N={N} ({len(rows)} sampled rows)", np.ascontiguousarray(got[:, rows]), w[rows], x)


@pytest.mark.parametrize("M,K,N", [(4096, 14336, 1), (14336, 4096, 8), (28672, 8192, 1), (8192, 28672, 8)])
def test_real_shape_integers_from_the_product_kernel(q4k, M, K, N):
    """The DUMP instantiation of whatever kernel the dispatch picks at this shape, on the sampled rows (as their own small matrix:
    the integer code does not depend on the grid) — and the fp32 bits of the sampled rows equal the full launch's bits."""
    from oracle import port
    w, x, rows = build_case(M, K, N, seed=M + K + N)
    rows = rows[:96]
    ws = np.ascontiguousarray(w[rows])
    rp, rsi, rsm = port.partials_matrix(ws, port.quantize_q8_K(x))
    dst, p, si, sm = q4k.mul_mat_dump(t(ws), t(x), q4k.DUMP_DISPATCH)
    assert np.array_equal(p.cpu().numpy(), rp) and np.array_equal(si.cpu().numpy(), rsi) and np.array_equal(sm.cpu().numpy(), rsm)
    full = q4k.mul_mat(t(w), t(x)).cpu().numpy()[:, rows]
    big = M * (K // 256) * 144 >= (34 << 20) and M >= 8192 and N == 1     # the large-matrix kernel: the sample alone runs the tensor-pipe kernel, same bits
    assert np.array_equal(dst.cpu().numpy().view(np.uint32), full.view(np.uint32)) or big
    if big:
        lib = q4k.load()
        try:
            lib.q4k_debug_set_big_matrix_bytes(1)
            dst2, p2, si2, sm2 = q4k.mul_mat_dump(t(ws), t(x), q4k.DUMP_BIG)
        finally:
            lib.q4k_debug_set_big_matrix_bytes(34 << 20)
        assert np.array_equal(p2.cpu().numpy(), rp) and np.array_equal(si2.cpu().numpy(), rsi) and np.array_equal(sm2.cpu().numpy(), rsm)
        assert np.array_equal(dst2.cpu().numpy().view(np.uint32), full.view(np.uint32))
