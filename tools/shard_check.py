"""torchrun --nproc-per-node G tools/shard_check.py : the row-sharded mul_mat on G GPUs (all three exchanges: NCCL all-gather,
round-1 peer stores + flag kernel, LL pairs with the in-kernel gather) against (a) the same matrix computed whole on every rank
(bit-equality expected: rows are independent and the kernel is the same) and (b) the CPU oracle on sampled rows.  The LL exchange
is also replayed from a CUDA graph with FRESH activations every replay (stale pairs of an earlier replay must never pass as
current), and exercised with ranks that own no rows."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402

# the row-sharded path runs the decode kernel (gemv_q4k_mma.cu / gemv_q4k_tma.cu) for every N <= 8; the single-GPU comparison must
# use the same kernel family (q4k_mul_mat alone hands N >= 3 to the batched decode kernel, whose float association differs)
q4k_b200.debug_set("wide_min", 9)
from q4k_b200.shard import RowShardedQ4K, shard_rows  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True


def oracle_rows(w, x, got, k=48):
    """rank 0: k sampled rows of the full matrix against the oracle restatement (scaled 1e-5 bound of tests/parity_util.py)"""
    from oracle import port
    M = w.shape[0]
    rows = np.unique(np.concatenate([np.arange(0, min(16, M)), np.arange(max(0, M // 2 - 8), min(M, M // 2 + 8)), np.arange(max(0, M - 16), M)]))[:k]
    wr, xn = w[torch.from_numpy(rows).to(w.device)].cpu().numpy(), x.cpu().numpy()
    ref, f64 = port.mul_mat(wr, xn, port.ORDER_AVX2), port.mul_mat_f64(wr, xn)
    rms = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
    g = got[:, torch.from_numpy(rows).to(got.device)].cpu().numpy()
    return bool(np.all(np.abs(g.astype(np.float64) - ref) <= 1e-5 * np.maximum(np.abs(ref), rms)))


for (M, K, N) in [(28672, 8192, 1), (8192, 28672, 1), (4100, 2048, 3), (2048, 2048, 8), (12, 2048, 2), (4, 2048, 2), (3584, 8192, 1)]:
    g = torch.Generator(device=dev); g.manual_seed(11)
    nb = K // 256
    w = torch.randint(0, 256, (M, nb, 144), dtype=torch.uint8, device=dev, generator=g)
    w[..., 0] = 0; w[..., 1] = 0x2c; w[..., 2] = 0; w[..., 3] = 0x28
    w = w.view(M, nb * 144)
    x = torch.randn((N, K), device=dev, generator=g)
    want = q4k_b200.mul_mat(w, x)
    r0, r1 = shard_rows(M, world)[rank]
    for gather in ("nccl", "peer", "ll"):
        sh = RowShardedQ4K(w[r0:r1].contiguous(), M, K, gather=gather)
        for _ in range(40 if M <= 4100 else 3):        # a burst of back-to-back calls: ranks that own no rows (M = 4, 12) must hold the others back
            got = sh.mul_mat(x).clone()
        torch.cuda.synchronize()
        same = bool(torch.equal(got.view(torch.int32), want.view(torch.int32)))
        orc = oracle_rows(w, x, got) if rank == 0 else True
        replay = True
        if gather == "ll":
            # graph replay with fresh activations each time: the kernel's epoch / buffer half come from the device-side counter
            xs = torch.empty_like(x)
            xs.copy_(x)
            out = torch.empty((N, M), device=dev)
            sh.mul_mat(xs)
            torch.cuda.synchronize()
            cg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(cg):
                out.copy_(sh.mul_mat(xs))
                out.copy_(sh.mul_mat(xs))
                out.copy_(sh.mul_mat(xs))          # an ODD number of calls per replay on purpose
            for it in range(5):
                xi = torch.randn((N, K), device=dev, generator=g)
                xs.copy_(xi)
                cg.replay()
                torch.cuda.synchronize()
                replay &= bool(torch.equal(out.view(torch.int32), q4k_b200.mul_mat(w, xi).view(torch.int32)))
        ok &= same and orc and replay
        if rank == 0:
            print(f"M={M} K={K} N={N} requested={gather} used={sh.gather} bit_equal_to_1gpu={same} oracle_rows_ok={orc} graph_replay_ok={replay} {sh.peer_note}", flush=True)
        dist.barrier()
t = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("SHARD_CHECK", "OK" if int(t) == 1 else "FAILED", flush=True)
dist.destroy_process_group()
os._exit(0 if int(t) == 1 else 1)
