"""GPU: the tensor-pipe decode kernel (gemv_q4k_mma.cu; float order "fast") against the oracle.  The integer stage must
be exact (checked with weights / activations whose float stage is exact in any order); fp32 within 1e-5 (assert_close of
test_gpu_parity.py: |got - ref| <= 1e-5 * max(|ref|, column rms))."""
import numpy as np
import pytest

from conftest import make_acts, make_weights
from parity_util import check
from test_gpu_parity import assert_close, t

pytestmark = pytest.mark.gpu


@pytest.fixture()
def fast_order(q4k):
    lib = q4k.load()
    lib.q4k_debug_set_gemv_order(2)
    q4k.debug_set("wide_min", 9)            # N = 3..8 through THIS kernel (q4k_mul_mat hands them to gemv_q4k_wide.cu by default)
    yield lib
    lib.q4k_debug_set_gemv_order(0)
    q4k.debug_set("wide_min", 2)


@pytest.mark.parametrize("M,K,N", [(5632, 2048, 1), (2048, 5632, 1), (100, 512, 1), (1, 256, 1), (17, 2560, 1), (64, 14336, 1),
                                   (257, 2048, 2), (300, 2048, 3), (129, 1024, 4), (77, 2048, 5), (40, 512, 6), (50, 768, 7),
                                   (64, 4096, 8), (33, 5632, 8), (32, 28672, 8), (20, 28672, 5)])
def test_mma_gemv_within_tolerance(q4k, fast_order, M, K, N):
    from oracle import port
    w, x = make_weights(M, K), make_acts(N, K)
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    check(f"tensor-pipe decode {M}x{K} N={N}", got, w, x)


@pytest.mark.parametrize("N", [1, 2, 5, 8])
def test_mma_gemv_integer_stage_is_exact(q4k, fast_order, N):
    """d = dmin = 1 in every weight block, activations that quantize to themselves (d_x = 1), small scales: every float
    operation of the kernel is exact, so the output is sumi - summ bit for bit whatever the accumulation order."""
    from oracle import port
    rng = np.random.default_rng(7 + N)
    M, K = 100, 2560                                        # ragged tile (100 % 16) and ragged round (10 superblocks)
    w = rng.integers(0, 256, (M, K // 256, 144), dtype=np.uint8)
    w[..., 0:2] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w[..., 2:4] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w[..., 4:16] &= 0x33                                    # 6-bit scales / mins <= 3 (+ high bits <= 3 << 4): sums stay < 2^24
    w = w.reshape(M, -1)
    x = rng.integers(-20, 21, (N, K)).astype(np.float32)
    x[:, ::256] = -127.0                                    # block max -> iscale = 1, d = 1
    xq = port.quantize_q8_K(x)
    want = np.zeros((N, M), np.int64)
    for n in range(N):
        for m in range(M):
            _, si, sm = port.vec_dot_partials(w[m], xq[n])
            want[n, m] = int(si.sum()) - int(sm.sum())
    assert np.abs(want).max() < 2 ** 24
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    assert np.array_equal(got, want.astype(np.float32))


@pytest.mark.parametrize("N", [1, 4])
def test_mma_result_independent_of_launch_shape(q4k, fast_order, N):
    """the same rows give the same bits alone, inside a fused group launch, and as the tail of a taller matrix"""
    K = 2048
    wa, wb = make_weights(96, K, seed=3), make_weights(40, K, seed=4)
    x = t(make_acts(N, K))
    a, b = q4k.mul_mat(t(wa), x), q4k.mul_mat(t(wb), x)
    ga, gb = q4k.mul_mat_group([t(wa), t(wb)], [x, x])
    tall = q4k.mul_mat(t(np.concatenate([wa, wb])), x)
    for u, v in ((a, ga), (b, gb), (a, tall[:, :96]), (b, tall[:, 96:])):
        assert np.array_equal(u.cpu().numpy().view(np.uint32), v.cpu().numpy().view(np.uint32))


def test_mma_q8_input_equals_fused_f32(q4k, fast_order):
    M, K, N = 200, 1024, 3
    w, x = t(make_weights(M, K)), t(make_acts(N, K))
    xq = q4k.quantize_q8_K(x)
    assert np.array_equal(q4k.mul_mat_q8(w, xq).cpu().numpy().view(np.uint32), q4k.mul_mat(w, x).cpu().numpy().view(np.uint32))


@pytest.mark.parametrize("M,K,N", [(2048, 16384, 1), (100, 14336, 1), (48, 28672, 2), (300, 24576, 5), (33, 8448, 8)])
def test_mma_split_k_does_not_change_the_bits(q4k, fast_order, M, K, N):
    """K split over the CTAs of a cluster (1, 2 or 4 parts): the segment sums are added in the same order wherever they
    were computed, so every split gives the same bits — and all of them are within tolerance of the oracle."""
    from oracle import port
    w, x = make_weights(M, K), make_acts(N, K)
    wd, xd = t(w), t(x)
    outs = {}
    try:
        for kp in (1, 2, 4):
            q4k.debug_set("mma_kparts", kp)
            outs[kp] = q4k.mul_mat(wd, xd).cpu().numpy()
    finally:
        q4k.debug_set("mma_kparts", 0)
    policy = q4k.mul_mat(wd, xd).cpu().numpy()
    check(f"tensor-pipe decode split-K {M}x{K} N={N}", outs[1], w, x)
    for kp in (2, 4):
        assert np.array_equal(outs[kp].view(np.uint32), outs[1].view(np.uint32)), kp
    assert np.array_equal(policy.view(np.uint32), outs[1].view(np.uint32))


@pytest.mark.parametrize("M,K", [(300, 2048), (64, 5632), (33, 2560), (16, 14336), (9, 28672), (5, 256), (130, 8192), (40, 8448)])
def test_big_matrix_kernel_gives_the_tensor_pipe_kernels_bits(q4k, M, K):
    """N == 1 launches above q4k_debug_set_big_matrix_bytes run the lane-per-superblock bulk-copy kernel in the tensor-pipe
    kernel's float association (gemv_q4k_tma.cu ORD 2): same bits, so a row does not depend on the size of its matrix.
    Shapes cover one / several 32-superblock segments, ragged rounds (K/256 % 8 != 0) and ragged row groups."""
    import torch
    from oracle import port
    lib = q4k.load()
    w, x = make_weights(M, K, seed=M), make_acts(1, K)
    wd, xd = t(w), t(x)
    xq = q4k.quantize_q8_K(xd)
    try:
        lib.q4k_debug_set_big_matrix_bytes(0)                 # never: the tensor-pipe kernel
        a = q4k.mul_mat(wd, xd).cpu().numpy()
        aq = q4k.mul_mat_q8(wd, xq).cpu().numpy()
        lib.q4k_debug_set_big_matrix_bytes(1)                 # always: the bulk-copy kernel
        n0 = q4k.kernel_launches()
        b = q4k.mul_mat(wd, xd).cpu().numpy()
        assert q4k.kernel_launches() - n0 in (1, 2)           # 2: activations above the fuse limit are quantized by their own kernel
        bq = q4k.mul_mat_q8(wd, xq).cpu().numpy()
    finally:
        lib.q4k_debug_set_big_matrix_bytes(34 << 20)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert np.array_equal(a.view(np.uint32), aq.view(np.uint32)) and np.array_equal(a.view(np.uint32), bq.view(np.uint32))
    check(f"large-matrix decode {M}x{K}", b, w, x)


@pytest.mark.parametrize("big", [0, 1])
def test_fused_peer_sync_single_rank(q4k, big):
    """q4k_mul_mat_peers_sync with world = 1: the decode kernel carries the cross-GPU step (last CTA: epoch bump, flag
    publish / wait over zero peers).  Results equal q4k_mul_mat bit for bit, the epoch counts the calls, the CTA arrival
    counter is back at zero — on both decode kernels (big = 1 forces the bulk-copy kernel).  The real exchange runs in
    tools/shard_check.py on 2 / 4 / 8 GPUs."""
    import ctypes as C
    import torch
    lib = q4k.load()
    M, K = 300, 2048
    w, x = t(make_weights(M, K, seed=3)), t(make_acts(1, K))
    want = q4k.mul_mat(w, x)
    out = torch.zeros((1, M), device="cuda")
    flags = torch.zeros(64, dtype=torch.int32, device="cuda")
    epoch = torch.zeros(1, dtype=torch.int32, device="cuda")
    done = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.empty(max(q4k.workspace_bytes(K, 1), 16), dtype=torch.uint8, device="cuda")
    peers = (C.c_void_p * 1)(out.data_ptr())
    farr = (C.c_void_p * 1)(flags.data_ptr())
    st = torch.cuda.current_stream().cuda_stream
    try:
        lib.q4k_debug_set_big_matrix_bytes(1 if big else 0)
        for i in range(3):
            out.zero_()
            n0 = q4k.kernel_launches()
            rc = lib.q4k_mul_mat_peers_sync(w.data_ptr(), M, K, w.stride(0), x.data_ptr(), 1, K * 4, peers, 1, 0, M * 4,
                                            ws.data_ptr(), ws.numel(), farr, 1, 0, epoch.data_ptr(), done.data_ptr(), st)
            assert rc == 0
            assert q4k.kernel_launches() == n0 + 1                      # one launch: no separate signal kernel
            torch.cuda.synchronize()
            assert torch.equal(out.view(torch.int32), want.view(torch.int32))
            assert int(epoch.item()) == i + 1 and int(done.item()) == 0
    finally:
        lib.q4k_debug_set_big_matrix_bytes(34 << 20)
