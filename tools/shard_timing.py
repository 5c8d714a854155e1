"""torchrun --nproc-per-node G tools/shard_timing.py : where the time of a row-sharded decode mul_mat goes.  Per shape:
(a) this rank's slice alone (plain q4k_mul_mat), (b) q4k_mul_mat_peers without the cross-GPU step (timing only: results may
be read early), (c) peers + q4k_peer_signal_wait.  CUDA graphs of 24 calls over rotating weight copies, max over ranks."""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
import q4k_b200  # noqa: E402
from q4k_b200.shard import shard_rows  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
import torch.distributed._symmetric_memory as symm  # noqa: E402
lib = q4k_b200.load()
if os.environ.get("BIG_BYTES"):            # A/B: threshold of the big-matrix kernel (0 = always the tensor-pipe kernel)
    lib.q4k_debug_set_big_matrix_bytes(int(os.environ["BIG_BYTES"]))
gname = dist.group.WORLD.group_name


def timed(fn, calls=24, reps=10):
    fn(0); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(calls):
            fn(i)
    g.replay(); torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (calls * reps)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


for (M, K) in [(28672, 8192), (8192, 28672)]:
    r0, r1 = shard_rows(M, world)[rank]
    m = r1 - r0
    nb = K // 256
    copies = max(2, (300 << 20) // (m * nb * 144) + 1)
    w = torch.randint(0, 256, (copies, m, nb, 144), dtype=torch.uint8, device=dev)
    w[..., 0] = 0; w[..., 1] = 0x2c; w[..., 2] = 0; w[..., 3] = 0x28
    w = w.view(copies, m, nb * 144)
    x = torch.randn((1, K), device=dev)
    out = symm.empty((2, 1, M), dtype=torch.float32, device=dev)
    hdl = symm.rendezvous(out, gname)
    flags = symm.empty((64,), dtype=torch.int32, device=dev)
    fh = symm.rendezvous(flags, gname)
    flags.zero_(); out.zero_()
    epoch = torch.zeros(1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize(); hdl.barrier(channel=0)
    order = [rank] + [r for r in range(world) if r != rank]
    half = M * 4
    parr = [(C.c_void_p * world)(*[hdl.buffer_ptrs[r] + b * half for r in order]) for b in range(2)]
    farr = (C.c_void_p * world)(*list(fh.buffer_ptrs))
    ws = torch.empty(max(q4k_b200.workspace_bytes(K, 1), 16), dtype=torch.uint8, device=dev)
    local_out = torch.empty((1, m), device=dev)

    def st():
        return torch.cuda.current_stream().cuda_stream

    def f_local(i):
        assert lib.q4k_mul_mat(w[i % copies].data_ptr(), m, K, nb * 144, x.data_ptr(), 1, K * 4, local_out.data_ptr(), m * 4, ws.data_ptr(), ws.numel(), st()) == 0

    def f_peers(i):
        assert lib.q4k_mul_mat_peers_ws(w[i % copies].data_ptr(), m, K, nb * 144, x.data_ptr(), 1, K * 4, parr[i & 1], world, r0, M * 4, ws.data_ptr(), ws.numel(), st()) == 0

    def f_one(i):          # same kernel, a single (local) destination
        assert lib.q4k_mul_mat_peers(w[i % copies].data_ptr(), m, K, nb * 144, x.data_ptr(), 1, K * 4, parr[i & 1], 1, r0, M * 4, st()) == 0

    def f_full(i):
        f_peers(i)
        assert lib.q4k_peer_signal_wait(farr, world, rank, epoch.data_ptr(), st()) == 0

    done = torch.zeros(1, dtype=torch.int32, device=dev)

    def f_fused(i):        # compute + peer stores + flag exchange in ONE launch
        assert lib.q4k_mul_mat_peers_sync(w[i % copies].data_ptr(), m, K, nb * 144, x.data_ptr(), 1, K * 4, parr[i & 1], world, r0, M * 4,
                                          ws.data_ptr(), ws.numel(), farr, world, rank, epoch.data_ptr(), done.data_ptr(), st()) == 0

    def f_sig(i):
        assert lib.q4k_peer_signal_wait(farr, world, rank, epoch.data_ptr(), st()) == 0

    res = dict(local=timed(f_local), peers_kernel_local_dst=timed(f_one), peers_no_sync=timed(f_peers), peers_and_signal=timed(f_full), fused_sync=timed(f_fused), signal_only=timed(f_sig))
    if rank == 0:
        print(f"M={M} K={K} world={world} rows/rank={m}: " + ", ".join(f"{k} {v:.2f} us" for k, v in res.items()), flush=True)
    dist.barrier()
dist.destroy_process_group()
os._exit(0)
