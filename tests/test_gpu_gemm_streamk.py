"""The prefill GEMM's persistent grid (csrc/gemm_q4k_tc.cu): launches with more output tiles than SMs cut the (tile, superblock)
sequence into one equal range per SM, so tiles are split between CTAs at arbitrary superblocks — later parts parked in the
workspace, the part with the first superblocks adds them in K order and stores.
Construction of tests/test_gpu_parity.py::test_gemm_integer_stage_is_exact: d = dmin = 1 in every weight block and activations
that quantize to themselves, so the fp32 output IS sumi - summ (< 2^24, exact under any float association): a lost, doubled or
misplaced superblock, a stale parked tile or a race on the flags shows as a wrong integer."""
import numpy as np
import pytest

from test_gpu_parity import t

pytestmark = pytest.mark.gpu


def exact_case(rng, M, K, N):
    w = rng.integers(0, 256, (M, K // 256, 144), dtype=np.uint8)
    w[..., 0:2] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    w[..., 2:4] = np.frombuffer(np.float16(1.0).tobytes(), np.uint8)
    x = rng.integers(-127, 128, (N, K)).astype(np.float32)
    x[:, ::256] = -127.0                                    # block max -> iscale = 1, d = 1
    return w.reshape(M, -1), x


def want_at(w, xq, pairs):
    from oracle import port
    out = []
    for n, m in pairs:
        _, si, sm = port.vec_dot_partials(w[m], xq[n])
        out.append(int(si.sum()) - int(sm.sum()))
    return np.array(out, np.int64)


def sample_pairs(rng, M, N, k):
    """(n, m) pairs that cover every 128 x 128 output tile corner region plus random ones"""
    ms = np.unique(np.r_[np.arange(0, M, 128), np.arange(127, M, 128), M - 1])
    ns = np.unique(np.r_[np.arange(0, N, 128), np.arange(63, N, 64), N - 1])
    pairs = [(int(n), int(m)) for n in ns for m in ms]
    pairs += [(int(rng.integers(N)), int(rng.integers(M))) for _ in range(k)]
    return pairs


@pytest.mark.parametrize("M,K,N", [(1280, 512, 1920), (2304, 1024, 1100), (19000, 256, 200), (300, 2560, 8000)])
def test_split_tiles_give_exact_integers(q4k, M, K, N):
    """150..324 tiles on 148 SMs: every CTA's range starts and ends inside a tile"""
    from oracle import port
    rng = np.random.default_rng(M + N)
    w, x = exact_case(rng, M, K, N)
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    again = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), again.view(np.uint32))
    xq = port.quantize_q8_K(x)
    pairs = sample_pairs(rng, M, N, 1500)
    want = want_at(w, xq, pairs)
    assert np.abs(want).max() < 2 ** 24
    g = np.array([got[n, m] for n, m in pairs])
    assert np.array_equal(g, want.astype(np.float32))
    # every element is an integer of plausible size (a tile that was never stored would hold the allocator's garbage)
    assert np.array_equal(got, np.rint(got)) and np.abs(got).max() < 2 ** 24


def test_grouped_launch_with_split_tiles(q4k):
    """three matrices in one launch, 14 column tiles: 168 tiles, ranges cross matrix boundaries"""
    from oracle import port
    rng = np.random.default_rng(3)
    K, N = 768, 1700
    Ms = (700, 650, 130)
    cases = [exact_case(rng, M, K, N) for M in Ms]
    x = cases[0][1]
    ws = [t(c[0]) for c in cases]
    n0 = q4k.kernel_launches()
    outs = q4k.mul_mat_group_ws(ws, t(x))
    assert q4k.kernel_launches() - n0 == 2
    xq = port.quantize_q8_K(x)
    for (w, _), o, M in zip(cases, outs, Ms):
        got = o.cpu().numpy()
        pairs = sample_pairs(rng, M, N, 600)
        g = np.array([got[n, m] for n, m in pairs])
        assert np.array_equal(g, want_at(w, xq, pairs).astype(np.float32))
        assert np.array_equal(got, np.rint(got))


def test_real_weights_through_split_tiles_stay_within_tolerance(q4k):
    from oracle import port
    from conftest import make_acts, make_weights
    from parity_util import check
    M, K, N = 1408, 1024, 1800                          # 11 x 15 = 165 tiles
    w, x = make_weights(M, K), make_acts(N, K)
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    rows = np.unique(np.r_[0:40, 600:680, M - 40:M])
    check("gemm stream-k", got[:, rows], w[rows], x)


@pytest.mark.parametrize("M,K,N", [(2048, 2048, 512), (1000, 768, 200), (300, 5632, 384)])
def test_half_tile_units_give_the_whole_tile_bits(q4k, M, K, N):
    """launches whose tiles fill at most half the SMs take 64-column half tiles as work units (twice the CTAs): every output still sums
    its whole K inside one CTA in K order, so the bits are those of the whole-tile grid (gemm_ksplit = 1 forces that) and the oracle
    tolerance holds"""
    from conftest import make_acts, make_weights
    from parity_util import check
    w, x = make_weights(M, K, seed=M), make_acts(N, K)
    got = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    try:
        q4k.debug_set("gemm_ksplit", 1)
        whole = q4k.mul_mat(t(w), t(x)).cpu().numpy()
    finally:
        q4k.debug_set("gemm_ksplit", 0)
    if K // 256 < 16:                    # (a long K is also cut between two CTAs per half tile: another association, tolerance only)
        assert np.array_equal(got.view(np.uint32), whole.view(np.uint32))
    rows = np.unique(np.r_[0:24, M // 2:M // 2 + 24, M - 24:M])
    check("gemm half tiles", got[:, rows], w[rows], x)
