"""GPU: horizontally fused launches (q4k_mul_mat_group) and graph plans (q4k_plan_*) give exactly the bytes the
one-launch-per-mul_mat path gives, which test_gpu_parity.py pins against the oracle; one case is also checked against
the oracle directly so this file stands on its own."""
import numpy as np
import pytest

from conftest import make_acts, make_weights

pytestmark = pytest.mark.gpu


def t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bits(a):
    return a.cpu().numpy().view(np.uint32)


@pytest.fixture(params=["default", "exact"], autouse=True)
def order(request, q4k):
    """every test here runs with both decode kernels: the default (tensor-pipe, tolerance order) and the AVX2-order one"""
    lib = q4k.load()
    lib.q4k_debug_set_gemv_order(1 if request.param == "exact" else 0)
    q4k.debug_set("wide_min", 9)        # N = 3..8 stay on the decode kernels here: q4k_mul_mat alone would hand them to gemv_q4k_wide.cu
    yield request.param
    lib.q4k_debug_set_gemv_order(0)
    q4k.debug_set("wide_min", 2)


@pytest.mark.parametrize("Ms,K,N", [((2048, 256, 256), 2048, 1), ((5632, 5632), 2048, 1), ((37, 4, 129, 1), 512, 2),
                                    ((64, 64), 4096, 8), ((300,), 1024, 3)])
def test_group_shared_input_equals_separate_launches(q4k, Ms, K, N):
    ws = [t(make_weights(M, K, seed=11 + i)) for i, M in enumerate(Ms)]
    x = t(make_acts(N, K))
    outs = q4k.mul_mat_group(ws, [x] * len(ws))
    for w, o in zip(ws, outs):
        assert np.array_equal(bits(o), bits(q4k.mul_mat(w, x)))


def test_group_distinct_inputs(q4k):
    K, N = 1024, 2
    ws = [t(make_weights(M, K, seed=3 + i)) for i, M in enumerate((96, 40, 200))]
    xs = [t(make_acts(N, K, seed=20 + i)) for i in range(3)]
    xs[2] = xs[0]                                            # a repeat shares the staged copy
    outs = q4k.mul_mat_group(ws, xs)
    for w, x, o in zip(ws, xs, outs):
        assert np.array_equal(bits(o), bits(q4k.mul_mat(w, x)))


def test_group_against_oracle(q4k, order):
    from oracle import port
    from test_gpu_parity import assert_close
    K = 2048
    wn = [make_weights(M, K, seed=5 + i) for i, M in enumerate((128, 36))]
    xn = make_acts(1, K)
    outs = q4k.mul_mat_group([t(w) for w in wn], [t(xn)] * 2)
    for w, o in zip(wn, outs):
        ref = port.mul_mat(w, xn, port.ORDER_AVX2)
        if order == "exact":
            assert np.array_equal(bits(o), ref.view(np.uint32))
        else:
            assert_close(o.cpu().numpy(), ref, port.mul_mat_f64(w, xn))


def test_group_rejects_bad_arguments(q4k):
    import torch
    L = q4k.load()
    assert L.q4k_mul_mat_group(0, None, None, 256, 144, None, 1, 1024, None, None) == -5
    assert L.q4k_mul_mat_group(5, None, None, 256, 144, None, 1, 1024, None, None) == -5
    assert L.q4k_mul_mat_group(1, None, None, 256, 144, None, 9, 1024, None, None) == -5     # N > 8: not a decode launch
    assert L.q4k_mul_mat_group(1, None, None, 100, 144, None, 1, 400, None, None) == -1
    del torch


def _layer(q4k, N, seed=0):
    """one transformer layer's Q4_K mul_mats the way the reference graph has them: q/k/v share an input, gate/up too"""
    K, FF = 512, 1024
    shapes = [("q", K, K, 0), ("k", 128, K, 0), ("v", 128, K, 0), ("o", K, K, 1), ("gate", FF, K, 2), ("up", FF, K, 2), ("down", K, FF, 3)]
    ws = [t(make_weights(M, Kk, seed=seed + i)) for i, (_, M, Kk, _) in enumerate(shapes)]
    return shapes, ws


@pytest.mark.parametrize("N", [1, 4, 16])
def test_plan_replay_equals_individual_mul_mats(q4k, N):
    import torch
    shapes, ws = _layer(q4k, N)
    plan = q4k.GraphPlan(0)
    idx = [plan.add(w, N, iid) for w, (_, _, _, iid) in zip(ws, shapes)]
    plan.finalize()
    assert idx == list(range(len(ws)))
    if N <= 8:
        assert plan.launches == 4                      # q+k+v, o, gate+up, down (quantizers fused in)
    xs = {iid: t(make_acts(N, Kk, seed=40 + iid)) for _, _, Kk, iid in shapes}
    for rep in range(2):                               # replays see fresh inputs
        for iid, x in xs.items():
            plan.input(iid).copy_(x + rep)
        n0 = q4k.kernel_launches()
        plan.run()
        torch.cuda.synchronize()
        assert q4k.kernel_launches() - n0 == plan.launches
        for i, (w, (_, _, _, iid)) in enumerate(zip(ws, shapes)):
            assert np.array_equal(bits(plan.output(i)), bits(q4k.mul_mat(w, xs[iid] + rep))), (rep, i)
    plan.close()


def test_plan_run_host_one_copy_each_way(q4k):
    import torch
    N = 1
    shapes, ws = _layer(q4k, N, seed=9)
    plan = q4k.GraphPlan(0)
    for w, (_, _, _, iid) in zip(ws, shapes):
        plan.add(w, N, iid)
    plan.finalize()
    hin = torch.zeros(plan.input_bytes, dtype=torch.uint8).pin_memory()
    hout = torch.zeros(plan.output_bytes, dtype=torch.uint8).pin_memory()
    xs = {}
    for _, _, Kk, iid in shapes:
        xs[iid] = make_acts(N, Kk, seed=70 + iid)
        off = plan.input_offset(iid)
        assert off >= 0 and off % 256 == 0
        hin[off: off + N * Kk * 4] = torch.from_numpy(xs[iid].view(np.uint8).reshape(-1))
    assert plan.input_offset(99) == -1 and plan.output_offset(len(ws)) == -1
    plan.run_host(hin, hout)
    torch.cuda.synchronize()
    for i, (w, (_, M, _, iid)) in enumerate(zip(ws, shapes)):
        off = plan.output_offset(i)
        got = hout[off: off + N * M * 4].numpy().view(np.uint32).reshape(N, M)
        assert np.array_equal(got, bits(q4k.mul_mat(w, t(xs[iid]))))
    plan.close()


def test_plan_run_host_pipeline_multi_layer(q4k):
    """Plans of >= 8 mul_mats replay from host buffers as a copy / compute pipeline (four graphs; the inputs of the later parts
    and every part's results move on a side stream).  Three layers, fresh inputs on every replay: every result equals the
    individual q4k_mul_mat bit for bit, also when replays are queued back to back without a synchronize in between."""
    import torch
    N, layers = 1, 3
    plan = q4k.GraphPlan(0)
    items = []                                            # (weights, global input id, K, M)
    for il in range(layers):
        shapes, ws = _layer(q4k, N, seed=100 * il)
        for w, (_, M, Kk, iid) in zip(ws, shapes):
            gid = 4 * il + iid
            plan.add(w, N, gid)
            items.append((w, gid, Kk, M))
    plan.finalize()
    hin = [torch.zeros(plan.input_bytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
    hout = [torch.zeros(plan.output_bytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
    xs = [{}, {}]
    for rep in range(2):
        for _, gid, Kk, _ in items:
            if gid not in xs[rep]:
                xs[rep][gid] = make_acts(N, Kk, seed=500 + 37 * rep + gid)
                off = plan.input_offset(gid)
                hin[rep][off: off + N * Kk * 4] = torch.from_numpy(xs[rep][gid].view(np.uint8).reshape(-1))
    plan.run_host(hin[0], hout[0])
    plan.run_host(hin[1], hout[1])                        # queued behind the first replay: must not overwrite inputs in use
    torch.cuda.synchronize()
    for rep in range(2):
        for i, (w, gid, _, M) in enumerate(items):
            off = plan.output_offset(i)
            got = hout[rep][off: off + N * M * 4].numpy().view(np.uint32).reshape(N, M)
            assert np.array_equal(got, bits(q4k.mul_mat(w, t(xs[rep][gid])))), (rep, i)
    plan.close()


def test_plan_argument_errors(q4k):
    w = t(make_weights(8, 512))
    plan = q4k.GraphPlan(0)
    L = q4k.load()
    assert L.q4k_plan_finalize(plan._p) == -5                       # empty plan
    assert plan.add(w, 1, 0) == 0
    with pytest.raises(q4k.Q4KError):
        plan.add(w, 2, 0)                                           # same input_id, different N
    assert L.q4k_plan_run(plan._p, None) == -5                      # not finalized
    plan.finalize()
    with pytest.raises(q4k.Q4KError):
        plan.add(w, 1, 0)                                           # frozen
    assert L.q4k_plan_run_host(plan._p, None, None, None) == -2
    plan.close()


def test_plan_default_dispatch_batch_4(q4k, order):
    """default knobs at N = 4: q/k/v and gate/up go out as group launches of the decode kernel, o and down through the batched
    decode kernel (quantizer + kernel): 2 + 2 * 2 launches; every output within 1e-5 of the oracle."""
    import torch
    from parity_util import check
    if order == "exact":
        pytest.skip("the AVX2-order mode keeps N <= 8 on the decode kernel")
    q4k.debug_set("wide_min", 2)
    N = 4
    shapes, ws = _layer(q4k, N, seed=3)
    plan = q4k.GraphPlan(0)
    for w, (_, _, _, iid) in zip(ws, shapes):
        plan.add(w, N, iid)
    plan.finalize()
    assert plan.launches == 6
    xs = {iid: make_acts(N, Kk, seed=70 + iid) for _, _, Kk, iid in shapes}
    for iid, x in xs.items():
        plan.input(iid).copy_(t(x))
    plan.run()
    torch.cuda.synchronize()
    for i, (w, (name, _, _, iid)) in enumerate(zip(ws, shapes)):
        check(f"plan N=4 {name}", plan.output(i).cpu().numpy(), w.cpu().numpy(), xs[iid])
    plan.close()


@pytest.mark.parametrize("Ms,K,N", [((2048, 256, 256), 2048, 130), ((640, 640), 1024, 512), ((100, 37, 300, 5), 512, 64), ((96,), 512, 200)])
def test_prefill_group_is_one_gemm_launch_with_the_single_launch_bits(q4k, order, Ms, K, N):
    """attn_q / attn_k / attn_v (or ffn_gate / ffn_up) at prefill batch sizes: one quantizer pass + ONE GEMM launch whose grid covers the
    row tiles of all matrices (ragged last tiles of each included); every output carries the bits of its own q4k_mul_mat."""
    if order == "exact":
        pytest.skip("float-order modes concern the decode kernels")
    ws = [t(make_weights(M, K, seed=31 + i)) for i, M in enumerate(Ms)]
    x = t(make_acts(N, K, seed=8))
    n0 = q4k.kernel_launches()
    outs = q4k.mul_mat_group_ws(ws, x)
    assert q4k.kernel_launches() - n0 == 2
    for w, o in zip(ws, outs):
        assert np.array_equal(bits(o), bits(q4k.mul_mat(w, x)))


def test_group_ws_declines_the_batched_decode_range(q4k):
    ws = [t(make_weights(64, 512, seed=1)), t(make_weights(32, 512, seed=2))]
    with pytest.raises(q4k.Q4KError) as e:
        q4k.mul_mat_group_ws(ws, t(make_acts(16, 512)))
    assert e.value.rc == -5
    outs = q4k.mul_mat_group_ws(ws, t(make_acts(4, 512)))                 # decode sizes: the decode group launch
    assert len(outs) == 2 and outs[0].shape == (4, 64)
