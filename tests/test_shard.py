"""Row-sharding host logic (ggml-neon-opt_b200/q4k_b200/shard.py): partition properties on CPU, and the N > 1 path with
two gloo ranks — each rank holds the oracle's result for ITS rows only and the gathered matrix must equal the oracle's
full result bit for bit (the exchange moves floats, it never recomputes them)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, ROOT)

from q4k_b200.shard import all_gather_rows, assemble_gathered, shard_rows  # noqa: E402


@pytest.mark.parametrize("M,world,granule", [(28672, 8, 4), (8192, 8, 128), (5632, 3, 4), (10, 4, 4), (0, 2, 4), (130, 2, 128), (7, 8, 4)])
def test_shard_rows_partition(M, world, granule):
    r = shard_rows(M, world, granule)
    assert len(r) == world and r[0][0] == 0 and r[-1][1] == M
    for (a0, a1), (b0, b1) in zip(r, r[1:]):
        assert a1 == b0 and a0 <= a1
    for r0, r1 in r:
        assert r0 % granule == 0 or r0 == M
    widths = [r1 - r0 for r0, r1 in r]
    assert max(widths) - min(widths) <= granule + (granule if M % granule else 0)


def test_70b_shapes_split_evenly():
    assert {r1 - r0 for r0, r1 in shard_rows(28672, 8)} == {3584}
    assert {r1 - r0 for r0, r1 in shard_rows(8192, 8, 128)} == {1024}


def test_assemble_gathered_trims_padding():
    ranges = [(0, 4), (4, 6)]
    g = torch.zeros((2, 3, 4))
    g[0] = torch.arange(12.0).view(3, 4)
    g[1, :, :2] = 100 + torch.arange(6.0).view(3, 2)
    out = assemble_gathered(g, ranges, 6)
    assert out.shape == (3, 6) and torch.equal(out[:, :4], g[0]) and torch.equal(out[:, 4:], g[1, :, :2])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, M, K, N, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import port as ora
        rng = np.random.default_rng(3)
        w = ora.quantize_q4_K(rng.uniform(-1, 1, (M, K)).astype(np.float32))
        x = rng.standard_normal((N, K)).astype(np.float32)
        ranges = shard_rows(M, world)
        r0, r1 = ranges[rank]
        local = torch.from_numpy(ora.mul_mat(w[r0:r1], x, ora.ORDER_AVX2))            # this rank's rows only
        full = all_gather_rows(local, ranges, M)
        want = torch.from_numpy(ora.mul_mat(w, x, ora.ORDER_AVX2))
        q.put((rank, bool(torch.equal(full.view(torch.int32), want.view(torch.int32))), tuple(full.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("M,K,N", [(64, 256, 1), (70, 512, 1), (36, 256, 3)])
def test_two_rank_gather_gloo(M, K, N):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, M, K, N, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True, (N, M)), (1, True, (N, M))]
