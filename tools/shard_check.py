"""torchrun --nproc-per-node G tools/shard_check.py : row-sharded mul_mat on G GPUs (both exchanges) against the same
matrix computed whole on every rank.  Bit-equality is expected: rows are independent and the kernel is the same."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import q4k_b200  # noqa: E402
from q4k_b200.shard import RowShardedQ4K, shard_rows  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True
for (M, K, N) in [(28672, 8192, 1), (8192, 28672, 1), (4100, 2048, 3), (2048, 2048, 16)]:
    g = torch.Generator(device=dev); g.manual_seed(11)
    nb = K // 256
    w = torch.randint(0, 256, (M, nb, 144), dtype=torch.uint8, device=dev, generator=g)
    w[..., 0] = 0; w[..., 1] = 0x2c; w[..., 2] = 0; w[..., 3] = 0x28
    w = w.view(M, nb * 144)
    x = torch.randn((N, K), device=dev, generator=g)
    want = q4k_b200.mul_mat(w, x)
    r0, r1 = shard_rows(M, world)[rank]
    for gather in ("nccl", "auto"):
        sh = RowShardedQ4K(w[r0:r1].contiguous(), M, K, gather=gather)
        for _ in range(3):
            got = sh.mul_mat(x)
        torch.cuda.synchronize()
        same = bool(torch.equal(got.view(torch.int32), want.view(torch.int32)))
        ok &= same
        if rank == 0:
            print(f"M={M} K={K} N={N} requested={gather} used={sh.gather} bit_equal={same} {sh.peer_note}", flush=True)
        dist.barrier()
t = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("SHARD_CHECK", "OK" if int(t) == 1 else "FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(t) == 1 else 1)
