"""GPU: the streamed single-column decode kernel (gemv_q4k_n1s.cu) — the data path of the batched decode kernel under the arithmetic
of the tensor-pipe decode kernel.  Its results must be the tensor-pipe kernel's BIT FOR BIT (same superblock step, same float
association), so a row's bits do not depend on the size of its matrix; that kernel's integers and fp32 are pinned against the
oracle in tests/test_gpu_mma.py / test_gpu_parity.py, one case is checked against the oracle here as well."""
import numpy as np
import pytest

from conftest import make_acts, make_weights
from parity_util import check

pytestmark = pytest.mark.gpu


def t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def run(q4k, w, x, streamed):
    q4k.debug_set("n1s_min_bytes", 1 if streamed else 0)
    q4k.debug_set("n1s_min_k", 256 if streamed else 1 << 40)
    try:
        n0 = q4k.kernel_launches()
        out = q4k.mul_mat(w, x).cpu().numpy()
        return out, q4k.kernel_launches() - n0
    finally:
        q4k.debug_set("n1s_min_bytes", 0)
        q4k.debug_set("n1s_min_k", 14336)


@pytest.mark.parametrize("M,K", [(64, 512), (300, 2048), (37, 2304), (100, 8448), (50, 14336), (37, 28672), (1, 256), (16, 256), (33, 9216),
                                 (4800, 1024)])
def test_streamed_kernel_bits_equal_the_decode_kernel(q4k, M, K):
    """ragged tiles, odd tile counts, ragged rounds (K / 256 % 8), one and several segments (K > 8192), more passes than CTAs"""
    w, x = make_weights(M, K, seed=M + K), make_acts(1, K, seed=K)
    wd, xd = t(w), t(x)
    a, la = run(q4k, wd, xd, True)
    b, lb = run(q4k, wd, xd, False)
    assert la == 2 and lb >= 1                      # quantizer + streamed kernel
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_streamed_kernel_against_the_oracle_and_default_dispatch(q4k):
    """4096 x 14336 (long K, 256 tiles: the default dispatch takes the streamed kernel): sampled rows against the oracle; full output equals
    the decode kernel's bits"""
    from test_gpu_real_shapes import build_case
    M, K = 4096, 14336
    w, x, rows = build_case(M, K, 1, seed=77)
    wd, xd = t(w), t(x)
    n0 = q4k.kernel_launches()
    got = q4k.mul_mat(wd, xd).cpu().numpy()
    assert q4k.kernel_launches() - n0 == 2
    check(f"streamed N=1 {M}x{K} ({len(rows)} sampled rows)", np.ascontiguousarray(got[:, rows]), w[rows], x)
    other, _ = run(q4k, wd, xd, False)
    assert np.array_equal(got.view(np.uint32), other.view(np.uint32))


def test_streamed_kernel_strided_weights(q4k):
    import torch
    M, K = 70, 4096
    w, x = make_weights(M, K, seed=3), make_acts(1, K, seed=4)
    nbw = K // 256 * 144
    wbuf = torch.zeros((M, nbw + 288), dtype=torch.uint8, device="cuda")
    wbuf[:, :nbw] = t(w)
    lib = q4k.load()
    ws = torch.empty(q4k.workspace_bytes(K, 1), dtype=torch.uint8, device="cuda")
    out = torch.zeros((1, M), device="cuda")
    q4k.debug_set("n1s_min_bytes", 1)
    try:
        rc = lib.q4k_mul_mat(wbuf.data_ptr(), M, K, wbuf.stride(0), t(x).data_ptr(), 1, K * 4, out.data_ptr(), M * 4, ws.data_ptr(), ws.numel(),
                             torch.cuda.current_stream().cuda_stream)
    finally:
        q4k.debug_set("n1s_min_bytes", 0)
    assert rc == 0
    dense, _ = run(q4k, t(w), t(x), True)
    assert np.array_equal(out.cpu().numpy().view(np.uint32), dense.view(np.uint32))
