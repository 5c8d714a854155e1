"""GPU, collected last: q4k_mul_mat_noalloc — the decode mul_mat with its Q8_K scratch in a static device array inside the
library (for callers that cannot cudaMalloc; no counterpart in the reference).
Called through the C ABI; checked against the existing entry point q4k_mul_mat on the same inputs — bit-identical, which is
inside the 1e-5 tolerance the other tests use — on two calls in a row (the array keeps the previous call's contents), and
against the CPU oracle.  Passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_acts, make_weights

pytestmark = pytest.mark.gpu

SCRATCH = 1 << 20        # Q4K_NOALLOC_SCRATCH_BYTES


def t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def noalloc(q4k, w, x, out):
    """the C-ABI call itself; out is preallocated by the caller: the entry point allocates nothing"""
    import torch
    lib = q4k.load()
    lib.q4k_mul_mat_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
    M, (N, K) = w.shape[0], x.shape
    return lib.q4k_mul_mat_noalloc(w.data_ptr(), M, K, w.stride(0), x.data_ptr(), N, K * 4, out.data_ptr(), M * 4,
                                   torch.cuda.current_stream().cuda_stream)


# (64, 28672, 1) and (128, 4096, 8): activations above the fuse limit -> the standalone quantizer writes the static scratch and
# the decode kernel reads it; (256, 2048, 1): fused launch (scratch unused); (40, 8192, 5): batch with a ragged tile
@pytest.mark.parametrize("M,K,N", [(64, 28672, 1), (128, 4096, 8), (256, 2048, 1), (40, 8192, 5)])
def test_noalloc_matches_the_existing_entry_point_twice_in_a_row(q4k, M, K, N):
    import torch
    from parity_util import check
    w = make_weights(M, K, seed=M + N)
    xs = [make_acts(N, K, seed=1), make_acts(N, K, seed=2, adversarial=False) * 3.0]
    wd = t(w)
    for call, x in enumerate(xs):                           # two calls in a row on the shared array, different activations
        xd = t(x)
        want = q4k.mul_mat(wd, xd)                          # existing entry point, caller-allocated workspace
        out = torch.full((N, M), float("nan"), device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        n0 = q4k.kernel_launches()
        rc = noalloc(q4k, wd, xd, out)
        assert rc == 0, (rc, q4k.load().q4k_last_cuda_error())
        torch.cuda.synchronize()
        assert q4k.kernel_launches() > n0                   # our kernels ran
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before      # and torch's allocator was not touched
        got, ref = out.cpu().numpy(), want.cpu().numpy()
        assert np.isfinite(got).all()
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), f"call {call}: differs from q4k_mul_mat"
        check(f"q4k_mul_mat_noalloc {M}x{K} N={N} call {call}", got, w, x)


def test_noalloc_rejects_what_does_not_fit(q4k):
    import torch
    lib = q4k.load()
    w, x = t(make_weights(8, 256)), torch.zeros((9, 256), device="cuda")
    out = torch.zeros((9, 8), device="cuda")
    assert noalloc(q4k, w, x, out) == -6                    # N > 8: not a decode shape
    # N = 8, K = 131072: 8 * 512 * 292 B = 1.14 MiB > the 1 MiB array.  Rejected before any pointer is touched.
    lib.q4k_mul_mat_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
    assert 8 * 512 * 292 > SCRATCH
    assert lib.q4k_mul_mat_noalloc(w.data_ptr(), 8, 131072, 512 * 144, x.data_ptr(), 8, 131072 * 4, out.data_ptr(), 32, 0) == -6
    assert lib.q4k_mul_mat_noalloc(w.data_ptr(), 8, 100, 144, x.data_ptr(), 1, 400, out.data_ptr(), 32, 0) == -1      # K % 256
    assert noalloc(q4k, w[:0], x[:1], out) == 0             # empty: nothing to do
