"""Row-sharded Q4_K mul_mat across the GPUs of one node.

Output rows are independent — dst[m, n] reads weight row m and activation column n only
(L/ggml/src/ggml-cpu/ggml-cpu.c:1192-1195) — so the weight matrix is split by rows once, at upload time (the
plugin's set_tensor), activations are replicated, every GPU runs quantize + GEMV/GEMM on its own rows, and the F32
output slices are exchanged:

  * `gather="nccl"`  one all_gather of [N, M/G] floats per GPU over NVLink/NVSwitch (torch.distributed, NCCL).
  * `gather="ll"`    (default when symmetric memory is available) compute + all-gather + completion in ONE launch
                     (q4k_mul_mat_peers_ll): the kernel's epilogue stores every result as an 8-byte {value, epoch} pair into the
                     LL buffer of EVERY rank (one atomic NVLink store each, no fence), and the tail of the same launch polls its
                     own LL buffer until this call's pairs have arrived from everybody and writes them out as plain floats.
                     No flag kernel; the epoch and the buffer half come from a device-side call counter, so a CUDA graph
                     replays it with no constraint on how many calls were captured.
  * `gather="peer"`  (round 1, kept for A/B) no collective kernel at all: the GEMV epilogue stores every result straight into the output
                     buffer of EVERY peer through NVLink-mapped symmetric memory (q4k_mul_mat_peers), followed by
                     one flag exchange (q4k_peer_signal_wait: a one-warp kernel, system-scope release / acquire on
                     peer-mapped flags, device-side epoch so that it replays inside a CUDA graph).  The payload is
                     tiny (28672 x 4 B per token for the 70B FFN), so this removes the NCCL launch + protocol
                     latency that dominates the baseline.

What the reference has instead: ggml-cuda's --split-mode row slices src0 rows by tensor_split, peer-copies src1 to
each device and copies the dst slices back to the main device only (L/ggml/src/ggml-cuda/ggml-cuda.cu:774-787,
1435-1760) — a gather-to-one with cudaMemcpyPeerAsync, no NCCL.  K-sharding is not offered: a split-K reduction
would change the float accumulation order of the superblock chain (and need an all-reduce).

One process per GPU.  Everything here except `RowShardedQ4K.mul_mat` itself is device-agnostic and is exercised
with world_size-2 gloo on CPU tensors in tests/test_shard.py.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

QK_K, Q4K_BLOCK = 256, 144


def shard_rows(M: int, world: int, granule: int = 4):
    """Contiguous row ranges [(r0, r1)] * world: balanced to within one granule, every boundary a multiple of
    `granule` (4 = rows of one GEMV tile; use 128 to keep GEMM row tiles whole).  Later ranks may be empty."""
    if M < 0 or world < 1 or granule < 1:
        raise ValueError("bad arguments")
    units = (M + granule - 1) // granule
    out = []
    for r in range(world):
        u0, u1 = units * r // world, units * (r + 1) // world
        out.append((min(u0 * granule, M), min(u1 * granule, M)))
    return out


def assemble_gathered(gathered: torch.Tensor, ranges, M: int) -> torch.Tensor:
    """gathered: [world, N, m_pad] (rank-major, zero padded to the widest slice) -> [N, M] with M fastest (ggml dst)."""
    world, N, _ = gathered.shape
    out = gathered.new_empty((N, M))
    for r, (r0, r1) in enumerate(ranges):
        if r1 > r0:
            out[:, r0:r1] = gathered[r, :, : r1 - r0]
    return out


def all_gather_rows(local: torch.Tensor, ranges, M: int, group=None) -> torch.Tensor:
    """local: this rank's [N, m_local] slice -> the full [N, M] on every rank.  One collective; equal slices at N = 1
    (the decode case) are gathered in place with no copy."""
    world = dist.get_world_size(group)
    N = local.shape[0]
    widths = [r1 - r0 for r0, r1 in ranges]
    m_pad = max(widths)
    if N == 1 and all(w == m_pad for w in widths):
        out = local.new_empty((1, M))
        dist.all_gather_into_tensor(out.view(world, m_pad), local.reshape(1, m_pad).contiguous(), group=group)
        return out
    padded = local
    if local.shape[1] != m_pad:
        padded = local.new_zeros((N, m_pad))
        padded[:, : local.shape[1]] = local
    gathered = local.new_empty((world, N, m_pad))
    dist.all_gather_into_tensor(gathered.view(world * N, m_pad), padded.contiguous(), group=group)   # concatenation form
    return assemble_gathered(gathered, ranges, M)


class RowShardedQ4K:
    """One rank's row slice of a Q4_K matrix [M, K] + the exchange that rebuilds the full [N, M] result on every rank.

    w_local : uint8 CUDA tensor [m_local, K/256*144] — rows ranges[rank] of the full matrix.
    gather  : "nccl" (all_gather of the slices), "ll" (one launch: LL pairs into every rank's buffer + in-kernel gather; needs
              torch symmetric memory over NVLink P2P; N <= 8) or "peer" (round 1: plain peer stores + a flag kernel).
              "auto" tries ll and falls back to nccl.  The result of the ll / peer modes is a view of a buffer that the next
              call (ll) / the call after the next (peer) of this object overwrites."""

    def __init__(self, w_local: torch.Tensor, M: int, K: int, group=None, gather: str = "auto", max_n: int = 8, granule: int = 4):
        from . import lib as _lib
        self._lib = _lib
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.M, self.K = M, K
        self.ranges = shard_rows(M, self.world, granule)
        r0, r1 = self.ranges[self.rank]
        if tuple(w_local.shape) != (r1 - r0, K // QK_K * Q4K_BLOCK):
            raise ValueError(f"rank {self.rank}: expected a [{r1 - r0}, {K // QK_K * Q4K_BLOCK}] slice, got {tuple(w_local.shape)}")
        self.w = w_local
        self.r0, self.r1 = r0, r1
        self.max_n = max_n
        self.gather = "nccl"
        self.peer_note = ""
        self._ws = None
        # The peer exchange alternates between two result buffers, chosen on the HOST per call.  A CUDA graph bakes the choice in,
        # so a captured sequence must contain an EVEN number of calls of this object (then replays keep alternating and the
        # "buffer (i + 1) % 2 is free after the exchange of call i" argument of _init_peer holds across replays too).
        self.needs_even_capture = False
        if gather in ("auto", "ll") and self.world > 1:
            try:
                self._init_ll()
                self.gather = "ll"
            except Exception as e:                       # no P2P / symmetric memory on this box: NCCL does the exchange
                if gather == "ll":
                    raise
                self.peer_note = f"LL exchange unavailable ({type(e).__name__}: {e}); using nccl"
        elif gather == "peer" and self.world > 1:
            self._init_peer()
            self.gather = "peer"
            self.needs_even_capture = True

    @classmethod
    def from_full(cls, w_full: torch.Tensor, K: int, device, group=None, **kw):
        """Shard at upload time: w_full is the whole matrix in HOST memory; only this rank's rows go to the GPU."""
        M = w_full.shape[0]
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        r0, r1 = shard_rows(M, world, kw.get("granule", 4))[rank]
        return cls(w_full[r0:r1].contiguous().to(device), M, K, group=group, **kw)

    def _init_ll(self):
        import ctypes as C
        import torch.distributed._symmetric_memory as symm
        dev = self.w.device
        gname = dist.group.WORLD.group_name if self.group is None else self.group.group_name
        # LL buffer: [2 halves][max_n][M] pairs of {float value, uint32 epoch} = one int64 each, zero-initialised (epoch 0 = never)
        nbytes = self._lib.load().q4k_ll_buffer_bytes(self.M, self.max_n)                  # data pairs + presence pairs
        self._ll = symm.empty((nbytes // 8,), dtype=torch.int64, device=dev)
        self._ll_hdl = symm.rendezvous(self._ll, gname)
        self._ll.zero_()
        self._ll_arr = (C.c_void_p * self.world)(*list(self._ll_hdl.buffer_ptrs))       # indexed by rank
        self._cnt = torch.zeros(2, dtype=torch.int32, device=dev)                       # call counter, CTA arrivals
        self._dst = torch.zeros((self.max_n, self.M), dtype=torch.float32, device=dev)  # plain result, complete when the launch is
        torch.cuda.synchronize(dev)
        self._ll_hdl.barrier(channel=0)

    def _init_peer(self):
        import ctypes as C
        import torch.distributed._symmetric_memory as symm
        dev = self.w.device
        # two result buffers used alternately: when the barrier of call i completes, every rank has launched call i,
        # hence (stream order) finished reading the results of call i - 1 ... so buffer (i + 1) % 2 is free again
        self._out = symm.empty((2, self.max_n, self.M), dtype=torch.float32, device=dev)
        self._hdl = symm.rendezvous(self._out, dist.group.WORLD.group_name if self.group is None else self.group.group_name)
        ptrs = list(self._hdl.buffer_ptrs)
        # local buffer first, then the peers
        order = [self.rank] + [r for r in range(self.world) if r != self.rank]
        half = self.max_n * self.M * 4
        self._peer_arr = [(C.c_void_p * self.world)(*[ptrs[r] + b * half for r in order]) for b in range(2)]
        self._flip = 0
        self._out.zero_()
        # flags of the cross-GPU step: one uint32 per (rank, peer) in symmetric memory + a local call counter
        self._flags = symm.empty((64,), dtype=torch.int32, device=dev)
        self._fhdl = symm.rendezvous(self._flags, dist.group.WORLD.group_name if self.group is None else self.group.group_name)
        self._flag_arr = (C.c_void_p * self.world)(*list(self._fhdl.buffer_ptrs))
        self._flags.zero_()
        self._epoch = torch.zeros(1, dtype=torch.int32, device=dev)
        self._done = torch.zeros(1, dtype=torch.int32, device=dev)           # CTA arrival counter of the fused cross-GPU step
        # Q4K_SHARD_FUSED_SYNC=1: the flag exchange fused into the decode launch (q4k_mul_mat_peers_sync) instead of the one-warp
        # kernel of its own.  Bit-equal, but measured SLOWER on B200 (2 GPUs: 25.3 vs 20.9 us per call at 14336x8192 per rank,
        # profiles/r1_shard_timing_fused_2gpu.txt): a fence in front of every CTA's arrival waits for that CTA's NVLink
        # stores, while the kernel boundary + a programmatically pre-launched one-warp kernel costs 5 us in all.  Off by default.
        self.fused_sync = os.environ.get("Q4K_SHARD_FUSED_SYNC", "0") == "1"
        torch.cuda.synchronize(dev)
        self._hdl.barrier(channel=0)
        self._fhdl.barrier(channel=0)

    def mul_mat(self, x: torch.Tensor) -> torch.Tensor:
        """x: replicated F32 activations [N, K] (identical on every rank) -> [N, M] on every rank.  With the peer
        exchange the result is a view of a buffer that is overwritten two calls later."""
        N = x.shape[0]
        lib = self._lib.load()
        st = torch.cuda.current_stream().cuda_stream
        if self.gather == "ll" and N <= self.max_n:
            if self._ws is None:
                self._ws = torch.empty(max(self._lib.workspace_bytes(self.K, self.max_n), 16), dtype=torch.uint8, device=self.w.device)
            rc = lib.q4k_mul_mat_peers_ll(self.w.data_ptr(), self.r1 - self.r0, self.K, self.w.stride(0), x.data_ptr(), N, x.stride(0) * 4,
                                          self._ll_arr, self.world, self.rank, self.M, self.max_n, self.r0,
                                          self._dst.data_ptr(), self.M * 4, self._cnt.data_ptr(), self._ws.data_ptr(), self._ws.numel(), st)
            if rc != 0:
                raise self._lib.Q4KError("q4k_mul_mat_peers_ll", rc, lib.q4k_last_cuda_error())
            return self._dst[:N]
        if self.gather == "peer" and N <= self.max_n:
            b = self._flip
            self._flip ^= 1
            if self._ws is None or self._ws.numel() < self._lib.workspace_bytes(self.K, N):
                self._ws = torch.empty(max(self._lib.workspace_bytes(self.K, self.max_n), 16), dtype=torch.uint8, device=self.w.device)
            if self.fused_sync:
                # compute + all-gather (peer stores) + cross-GPU flag exchange in ONE launch
                rc = lib.q4k_mul_mat_peers_sync(self.w.data_ptr(), self.r1 - self.r0, self.K, self.w.stride(0), x.data_ptr(), N, x.stride(0) * 4,
                                                self._peer_arr[b], self.world, self.r0, self.M * 4, self._ws.data_ptr(), self._ws.numel(),
                                                self._flag_arr, self.world, self.rank, self._epoch.data_ptr(), self._done.data_ptr(), st)
                if rc != 0:
                    raise self._lib.Q4KError("q4k_mul_mat_peers_sync", rc, lib.q4k_last_cuda_error())
                return self._out[b, :N]
            if self.r1 > self.r0:
                rc = lib.q4k_mul_mat_peers_ws(self.w.data_ptr(), self.r1 - self.r0, self.K, self.w.stride(0), x.data_ptr(), N, x.stride(0) * 4,
                                              self._peer_arr[b], self.world, self.r0, self.M * 4, self._ws.data_ptr(), self._ws.numel(), st)
                if rc != 0:
                    raise self._lib.Q4KError("q4k_mul_mat_peers_ws", rc, lib.q4k_last_cuda_error())
            rc = lib.q4k_peer_signal_wait(self._flag_arr, self.world, self.rank, self._epoch.data_ptr(), st)   # every rank's stores have landed
            if rc != 0:
                raise self._lib.Q4KError("q4k_peer_signal_wait", rc, lib.q4k_last_cuda_error())
            return self._out[b, :N]
        local = self._lib.mul_mat(self.w, x) if self.r1 > self.r0 else x.new_empty((N, 0))
        return all_gather_rows(local, self.ranges, self.M, self.group)
