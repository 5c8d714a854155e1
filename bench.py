#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the Q4_K x Q8_K mul_mat hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

ONE workload at every N (so that the driver's 1 -> 8 GPU curve is real strong scaling):

  "llama3-70b-ffn" (BASELINE.json configs[4], the default at every N): the Llama-3-70B FFN pair — Q4_K 28672x8192 (ffn_up shape)
  and 8192x28672 (ffn_down shape), batch 1 — FFN70B_COPIES distinct copies of the pair = 18 mul_mats, 2.38 GB of Q4_K weights
  per step, identical at N = 1 / 2 / 4 / 8.  N = 1: the 18 mul_mats as one q4k_plan (CUDA graph).  N > 1 (torchrun): weight rows
  sharded over the N ranks, activations replicated, F32 output slices exchanged so every rank holds the full result
  (SURVEY.md §8e).  Fixed total work -> "scaling": "strong".  Each mul_mat = quantize_row_q8_K of its F32 activations + the
  GEMV, as ggml_compute_forward_mul_mat does (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401).

  "tinyllama-token" (configs[1]; reported at N = 1 as the `decode_token` object of the default line, or as the headline with
  --workload tinyllama-token): every GGML_TYPE_Q4_K mul_mat of ONE decode token of TinyLlama-1.1B Q4_K_M, batch 1, in llama
  graph order — 134 mul_mats, 477 167 616 B of Q4_K weights.

`value` = algorithmic bytes of the step (M*K/256*144 + N*K*4 + M*N*4 per mul_mat) / device time, inputs resident
in HBM.  `e2e` = the same through the host-buffer entries (pinned host activations in, host results out, copies inside the timed
region).  `--impl reference` times the UNMODIFIED reference ggml CPU code (oracle/_ref, built by oracle/Makefile) on the host
cores on the same workload.  oracle/ is used here only as that CPU baseline and as the CHECKER of sampled output rows before the
timed region (`parity_checked`), never on the measured GPU path.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))

L2_BYTES = 126 << 20
HBM_NOMINAL_GBS = 8000.0           # BASELINE.json's stated denominator ("~8 TB/s")
HBM_FALLBACK_GBS = 6650.0          # /opt/skills/guides/B200_PROFILING.md fallback if MEASURED_PEAKS.json is absent
FFN70B_COPIES = 9                  # 9 x (28672x8192 + 8192x28672) = 18 mul_mats: > 2.2 x L2 of distinct weights per GPU even at N = 8
# tcgen05.mma kind::i8 rate measured on this pool's B200 by tools/micro/umma_rate.cu (profiles/r1_micro_umma_rate.txt), MAC/clk/SM
UMMA_I8_MAC_PER_CLK_SM = {"ts_n64": 6150, "ts_n128": 7030, "ts_n256": 7565}
SM_COUNT = 148
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the workload's dominant kernel, from that kernel's own
# `ncu --set full` capture (null where the workload's kernel / shape has not been captured)
NCU_TRAFFIC = {
    "ffn70b": {"bytes": 135900000, "kernel": "mean of gemv_q4k_tma_kernel<1, 2> at 28672x8192 N=1 and gemv_q4k_n1s_kernel at 8192x28672 N=1 (the two launches of the workload)",
               "source": "profiles/r2_ncu_q4k_big_28672x8192.txt and profiles/r2_ncu_n1s_8192x28672.txt: 132.2 MB DRAM read + 3.7-4.0 MB write per launch "
                         "(the write figure is what ncu attributes to the launch at device level; the kernels store 115 / 33 KB) vs 132 268 032 B algorithmic"},
    "tinyllama-token": {"bytes": 6562304, "kernel": "gemv_q4k_mma_kernel<1, 0> at 5632x2048 N=1",
                        "source": "profiles/r1_ncu_gemv_mma_c1_5632x2048.txt: 6 562 304 B DRAM read + 0 B write per launch vs 6 518 784 B algorithmic"},
}
Q6K_LAYERS = {0, 1, 4, 7, 10, 13, 16, 19, 20, 21}   # use_more_bits(i, 22), L/src/llama-quant.cpp:185-187


# ------------------------------------------------------------------------------------------------ workloads
def tinyllama_token_matmuls():
    """(name, M, K) of every Q4_K mul_mat of one TinyLlama-1.1B Q4_K_M decode token, in graph order.
    n_embd 2048, n_head_kv 4 x 64 = 256, n_ff 5632, 22 layers; attn_v / ffn_down are Q6_K in Q6K_LAYERS
    (L/src/llama-quant.cpp:302-303, 358-364) and output.weight is Q6_K -> not on this path."""
    mats = []
    for il in range(22):
        mats.append((f"blk.{il}.attn_q", 2048, 2048))
        mats.append((f"blk.{il}.attn_k", 256, 2048))
        if il not in Q6K_LAYERS:
            mats.append((f"blk.{il}.attn_v", 256, 2048))
        mats.append((f"blk.{il}.attn_output", 2048, 2048))
        mats.append((f"blk.{il}.ffn_gate", 5632, 2048))
        mats.append((f"blk.{il}.ffn_up", 5632, 2048))
        if il not in Q6K_LAYERS:
            mats.append((f"blk.{il}.ffn_down", 2048, 5632))
    return mats


def tinyllama_full_token_matmuls():
    """(name, M, K, is_q6k) of EVERY weight mul_mat of one TinyLlama-1.1B Q4_K_M decode token in graph order: the 134 Q4_K
    ones plus attn_v / ffn_down of Q6K_LAYERS and output.weight in Q6_K (L/src/llama-quant.cpp:228-230, 302-303, 358-364)."""
    mats = []
    for il in range(22):
        q6 = il in Q6K_LAYERS
        mats += [(f"blk.{il}.attn_q", 2048, 2048, False), (f"blk.{il}.attn_k", 256, 2048, False), (f"blk.{il}.attn_v", 256, 2048, q6),
                 (f"blk.{il}.attn_output", 2048, 2048, False), (f"blk.{il}.ffn_gate", 5632, 2048, False),
                 (f"blk.{il}.ffn_up", 5632, 2048, False), (f"blk.{il}.ffn_down", 2048, 5632, q6)]
    mats.append(("output", 32000, 2048, True))
    return mats


def shared_input_ids(mats):
    """input id per mul_mat: attn_q / attn_k / attn_v of a layer read the same attn_norm output and ffn_gate / ffn_up the
    same ffn_norm output (L/src/models/llama.cpp build_attn / build_ffn inputs); every other mul_mat has its own input."""
    ids, nxt, cur = [], -1, None
    for name, _, _ in mats:
        blk, _, leaf = name.rpartition(".")
        fam = "qkv" if leaf in ("attn_q", "attn_k", "attn_v") else ("gu" if leaf in ("ffn_gate", "ffn_up") else leaf)
        if (blk, fam) != cur or fam not in ("qkv", "gu"):
            nxt, cur = nxt + 1, (blk, fam)
        ids.append(nxt)
    return ids


def ffn70b_matmuls(copies=FFN70B_COPIES):
    return [(f"{n}.{c}", M, K) for c in range(copies) for n, M, K in (("ffn_up_70b", 28672, 8192), ("ffn_down_70b", 8192, 28672))]


def workload_mats(name):
    return tinyllama_token_matmuls() if name == "tinyllama-token" else ffn70b_matmuls()


def workload_name(w):
    return {"tinyllama-token": "tinyllama-1.1b-q4_k_m-decode-token (configs[1]): 134 Q4_K mul_mats, batch 1",
            "ffn70b": f"llama3-70b-ffn (configs[4]): {FFN70B_COPIES} x (28672x8192 + 8192x28672) Q4_K, batch 1, rows sharded over the GPUs"}[w]


def config_obj(workload, N, n_mats, world):
    """Identical for both arms (the driver compares the two config objects)."""
    return {"workload": workload_name(workload), "batch": N, "mul_mats": n_mats,
            "parallelism": ("rowshard" if workload == "ffn70b" else "replica") + str(world)}


def algo_bytes(M, K, N):
    return M * (K // 256) * 144 + N * K * 4 + M * N * 4


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", p
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)", {}


def hbm_fracs(gbs, peak):
    return {"frac_measured": round(gbs / peak, 4), "frac_nominal_8TBs": round(gbs / HBM_NOMINAL_GBS, 4)}


def int8_peaks(sm_mhz):
    """Measured kind::i8 MMA rate (tools/micro/umma_rate.cu) at the SM clock of this run -> TOPS.  The prefill kernel keeps the
    reference's integers exact, which takes two int8 MMAs per weight: its ceiling is half the MMA rate."""
    mhz = sm_mhz or 1965.0
    out = {k: round(v * 2 * SM_COUNT * mhz * 1e6 / 1e12, 1) for k, v in UMMA_I8_MAC_PER_CLK_SM.items()}
    out["sm_mhz"] = mhz
    out["source"] = "tools/micro/umma_rate.cu on this pool's B200 (profiles/r1_micro_umma_rate.txt), MAC/clk/SM x 2 x 148 SMs x the run's SM clock"
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons: one query before, a 20 ms sampling loop DURING the measured region, one query after."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.thr = index, [], None, None

    def _once(self):
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=10).stdout
            for line in out.strip().splitlines():
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def start(self):
        self._once()
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thr = threading.Thread(target=self._read, daemon=True)
        self.thr.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is not None:
            time.sleep(0.05)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        self._once()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["nvidia-smi unavailable"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "sm_mhz_min": min(sm), "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ GPU arm
def rand_q4k(torch, M, K, device, gen):
    """Random VALID block_q4_K bytes (any byte pattern with finite fp16 d/dmin is a valid block):
    throughput-only weights, d = 2^-4, dmin = 2^-5.  Parity is tested elsewhere on really quantized weights."""
    nb = K // 256
    w = torch.randint(0, 256, (M, nb, 144), dtype=torch.uint8, device=device, generator=gen)
    w[..., 0] = 0x00; w[..., 1] = 0x2c; w[..., 2] = 0x00; w[..., 3] = 0x28
    return w.view(M, nb * 144)


def rand_q6k(torch, M, K, device, gen):
    """Random VALID block_q6_K bytes (any ql / qh / scales bytes are valid), d = 2^-8: throughput-only weights."""
    nb = K // 256
    w = torch.randint(0, 256, (M, nb, 210), dtype=torch.uint8, device=device, generator=gen)
    w[..., 208] = 0x00
    w[..., 209] = 0x1c
    return w.view(M, nb * 210)


class MatSet:
    """Device-resident weights + per-mul_mat activations / results / staging for one workload."""

    def __init__(self, torch, q4k, mats, N, device, seed=1234, row_range=None):
        self.torch, self.q4k, self.lib, self.N = torch, q4k, q4k.load(), N
        gen = torch.Generator(device=device)
        gen.manual_seed(seed)
        self.items = []
        ids = shared_input_ids(mats)
        for (name, M, K), iid in zip(mats, ids):
            r0, r1 = (0, M) if row_range is None else row_range(M)
            w = rand_q4k(torch, r1 - r0, K, device, gen)
            x = torch.randn((N, K), device=device, generator=gen)
            # prefill: mul_mats of one layer that read the same tensor (q/k/v, gate/up) get the same activations, so the second
            # and third reuse the quantized copy (q4k_mul_mat_reuse — what the plugin's graph_compute does)
            shared = N > 64 and self.items and self.items[-1]["iid"] == iid and self.items[-1]["K"] == K
            if shared:
                x = self.items[-1]["x"]
            dst = torch.zeros((N, r1 - r0), device=device)
            self.items.append(dict(name=name, M=r1 - r0, K=K, M_full=M, r0=r0, w=w, x=x, dst=dst, iid=iid, reuse=bool(shared)))
        kmax = max(K for _, _, K in mats)
        self.ws = torch.empty(max(q4k.workspace_bytes(kmax, N), 16), dtype=torch.uint8, device=device)
        self.algo_bytes = sum(algo_bytes(it["M"], it["K"], N) for it in self.items)
        self.weight_bytes = sum(it["w"].numel() for it in self.items)

    def enqueue_device(self):
        """One step with inputs resident in HBM: quantize + GEMV per mul_mat on the current stream."""
        st = self.torch.cuda.current_stream().cuda_stream
        ws, wsn, N, f = self.ws.data_ptr(), self.ws.numel(), self.N, self.lib.q4k_mul_mat
        import ctypes as C
        i = 0
        while i < len(self.items):
            it = self.items[i]
            j = i + 1
            while j < len(self.items) and j - i < 4 and self.items[j]["reuse"]:
                j += 1
            if j - i > 1:          # prefill mul_mats of a layer on one activation tensor (q/k/v, gate/up): one quantizer pass, one GEMM launch
                g = self.items[i:j]
                n = len(g)
                rc = self.lib.q4k_mul_mat_group_ws(n, (C.c_void_p * n)(*[x["w"].data_ptr() for x in g]), (C.c_int64 * n)(*[x["M"] for x in g]), it["K"],
                                                   it["w"].stride(0), it["x"].data_ptr(), N, it["K"] * 4,
                                                   (C.c_void_p * n)(*[x["dst"].data_ptr() for x in g]), ws, wsn, st)
                if rc != 0:
                    raise RuntimeError(f"q4k_mul_mat_group_ws rc={rc} cuda={self.lib.q4k_last_cuda_error()}")
                i = j
                continue
            i = j
            if it["reuse"]:
                rc = self.lib.q4k_mul_mat_reuse(it["w"].data_ptr(), it["M"], it["K"], it["w"].stride(0), N, it["dst"].data_ptr(), it["M"] * 4, ws, wsn, st)
            else:
                rc = f(it["w"].data_ptr(), it["M"], it["K"], it["w"].stride(0), it["x"].data_ptr(), N, it["K"] * 4,
                       it["dst"].data_ptr(), it["M"] * 4, ws, wsn, st)
            if rc != 0:
                raise RuntimeError(f"q4k_mul_mat rc={rc} cuda={self.lib.q4k_last_cuda_error()}")

    def make_plan(self):
        """The whole step as one q4k_plan: packed inputs / outputs, shared activations, fused launches, one CUDA graph."""
        t = self.torch
        ids = shared_input_ids([(it["name"], it["M"], it["K"]) for it in self.items])
        plan = self.q4k.GraphPlan(self.items[0]["w"].device.index)
        for it, iid in zip(self.items, ids):
            it["plan_index"], it["input_id"] = plan.add(it["w"], self.N, iid), iid
        plan.finalize()
        for it in self.items:
            plan.input(it["input_id"]).copy_(it["x"])
        self.plan = plan
        self.plan_in_host = t.zeros(plan.input_bytes, dtype=t.uint8).pin_memory()
        self.plan_out_host = t.zeros(plan.output_bytes, dtype=t.uint8).pin_memory()
        self.plan_in_host.copy_(plan._in)
        self.h2d, self.d2h = plan.input_bytes, plan.output_bytes
        return plan


def timed_steps(torch, step_fn, steps, warmup, barrier):
    """W untimed warm-ups, then exactly K steps between CUDA events on the launching stream, a barrier +
    synchronize on both sides.  Returns milliseconds for the K steps."""
    for _ in range(warmup):
        step_fn()
    torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step_fn()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    return e0.elapsed_time(e1)


def timed_region(torch, region_fn, warm_fn, barrier):
    """warm_fn() untimed, then region_fn() once between CUDA events, barrier + synchronize on both sides.  Milliseconds."""
    warm_fn()
    torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    region_fn()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    return e0.elapsed_time(e1)


def sustained(torch, step_fn, seconds, per_step_ms, barrier):
    """The same step repeated for ~`seconds` (device-timed): evidence that the K-step figure is not a burst-clock artefact."""
    n = max(10, int(seconds * 1e3 / max(per_step_ms, 1e-3)))
    ms = timed_steps(torch, step_fn, n, 0, barrier)
    return n, ms


def graph_of(torch, enqueue):
    """Capture one step into a CUDA graph (the launches are ours; the graph removes host launch latency)."""
    enqueue()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        enqueue()
    return g


def oracle_check_rows(w_rows, x, got_rows, what):
    """CHECKER (not measured): sampled output rows against the oracle restatement (AVX2 order), scaled 1e-5 bound of
    tests/parity_util.py plus the literal relative statistic.  w_rows: uint8 [R, nb*144] numpy; x: [N, K]; got_rows: [N, R]."""
    import numpy as np
    from oracle import port
    ref = port.mul_mat(w_rows, x, port.ORDER_AVX2)
    f64 = port.mul_mat_f64(w_rows, x)
    rms = np.sqrt((f64 ** 2).mean(axis=1, keepdims=True))
    with np.errstate(divide="ignore", invalid="ignore"):
        rel = np.abs(got_rows.astype(np.float64) - ref) / np.abs(ref)
    rel[got_rows == ref] = 0.0
    ok = bool(np.all(np.abs(got_rows.astype(np.float64) - ref) <= 1e-5 * np.maximum(np.abs(ref), rms)))
    return {"what": what, "rows": int(w_rows.shape[0]), "ok": ok, "max_rel": float(rel.max()), "n_above_1e-5_rel": int((rel > 1e-5).sum())}


def sample_positions(m, k):
    """k row positions spread over the first / middle / last rows of [0, m)."""
    import numpy as np
    k = min(k, m)
    third = max(1, k // 3)
    pos = list(range(0, third)) + list(range(m // 2, m // 2 + third)) + list(range(m - (k - 2 * third), m))
    return np.array(sorted(set(min(max(p, 0), m - 1) for p in pos)))


def parity_before_timing(torch, items, results, N, world, rank, dist, rows_per_matrix=64):
    """Every rank contributes the weights of a few of ITS rows of every matrix; rank 0 compares the corresponding rows of the
    (gathered) results it holds with the oracle.  Returns (all_ok, per-matrix summaries) on rank 0, (None, None) elsewhere."""
    import numpy as np
    per_rank = -(-rows_per_matrix // world)
    summaries, all_ok = [], True
    for it, full in zip(items, results):
        pos = sample_positions(it["M"], per_rank)
        wl = it["w"][torch.from_numpy(pos).to(it["w"].device)].contiguous()            # [R, nb*144] this rank's sampled rows
        gpos = torch.from_numpy(pos + it["r0"]).to(it["w"].device)
        if world > 1:
            R = torch.tensor([wl.shape[0]], device=wl.device)
            Rmax = R.clone()
            dist.all_reduce(Rmax, op=dist.ReduceOp.MAX)
            pad_w = torch.zeros((int(Rmax), wl.shape[1]), dtype=torch.uint8, device=wl.device)
            pad_w[: wl.shape[0]] = wl
            pad_p = torch.full((int(Rmax),), -1, dtype=torch.int64, device=wl.device)
            pad_p[: wl.shape[0]] = gpos
            ws = [torch.empty_like(pad_w) for _ in range(world)]
            ps = [torch.empty_like(pad_p) for _ in range(world)]
            dist.all_gather(ws, pad_w)
            dist.all_gather(ps, pad_p)
            wl, gpos = torch.cat(ws), torch.cat(ps)
            keep = gpos >= 0
            wl, gpos = wl[keep], gpos[keep]
        if rank == 0:
            got = full[:, gpos].cpu().numpy()
            s = oracle_check_rows(wl.cpu().numpy(), it["x"].cpu().numpy(), np.ascontiguousarray(got), it["name"])
            all_ok = all_ok and s["ok"]
            summaries.append(s)
    if rank != 0:
        return None, None
    worst = max((s["max_rel"] for s in summaries), default=0.0)
    return all_ok, {"matrices": len(summaries), "rows_per_matrix": summaries[0]["rows"] if summaries else 0, "all_within_scaled_1e-5": all_ok,
                    "max_rel": worst, "n_above_1e-5_rel": sum(s["n_above_1e-5_rel"] for s in summaries),
                    "checker": "oracle.port AVX2 order (bit-identical to oracle/_ref), tests/parity_util.py bound"}


# ------------------------------------------------------------------------------------------------ CPU figures (reference code)
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_figures_for_shape(M, K, N, unit, budget_macs=3e8):
    """BASELINE.md §5 timings of the UNMODIFIED reference (oracle/_ref) for one shape, each on a bounded row sample:
    (i) direct 1 thread (quantize_row_q8_K + ggml_vec_dot_q4_K_q8_K loop), (ii) whole op on ggml_backend_cpu with 1 thread and
    with all cores (plain CPU buffer -> vec_dot path), (iii) whole op, all cores, CPU_REPACK buffer (llama-bench's default on AVX2).
    unit "GB/s": algorithmic bytes of the sample / time; "TOPS": 2*m*K*N / time."""
    from oracle import ref
    threads = host_threads()

    def conv(m, sec):
        return round(algo_bytes(m, K, N) / sec / 1e9, 3) if unit == "GB/s" else round(2.0 * m * K * N / sec / 1e12, 4)

    m1 = int(min(M, max(64, budget_macs // (K * N)))) // 8 * 8 or 8
    mall = int(min(M, max(256, 16 * budget_macs // (K * N)))) // 8 * 8 or 8
    out = {"unit": unit, "threads_all": threads, "rows_1t": m1, "rows_all": mall}
    try:
        out["direct_1t"] = conv(m1, ref.direct_q4k(m1, K, N, 1, 3)[0])
        out["whole_op_1t"] = conv(m1, ref.mul_mat_set_ex([(m1, K, 0)], N, 1, 1, 3, False)[0])
        out["whole_op_all"] = conv(mall, ref.mul_mat_set_ex([(mall, K, 0)], N, threads, 1, 3, False)[0])
        out["whole_op_all_repack"] = conv(mall, ref.mul_mat_set_ex([(mall, K, 0)], N, threads, 1, 3, True)[0])
    except Exception as e:
        out["unavailable"] = str(e)
    return out


def cpu_baseline_obj(mats, N, total_bytes, budget_s=12.0, q6=None):
    """Reference CPU code on this box's host cores on the WHOLE step, bounded to ~budget_s per path: plain CPU buffer
    (ggml_compute_forward_mul_mat -> AVX2 vec_dot) and CPU_REPACK (what llama-bench runs by default on an AVX2 host)."""
    from oracle import ref
    if not ref.available():
        raise RuntimeError("oracle/_ref is not built (make -C oracle ref in the build container)")
    threads = host_threads()
    shapes = [(M, K, bool(q6[i]) if q6 else False) for i, (_, M, K) in enumerate(mats)]
    res = {}
    for path, rp in (("plain", False), ("repack", True)):
        try:
            t0 = time.time()
            best, mean, _ = ref.mul_mat_set_ex(shapes, N, threads, 1, 2, rp)
            iters = int(max(2, min(20, (budget_s - (time.time() - t0)) / max(mean, 1e-4))))
            best2, mean2, _ = ref.mul_mat_set_ex(shapes, N, threads, 1, iters, rp)
            mean, best = min(mean, mean2), min(best, best2)
            res[path] = {"value": round(total_bytes / mean / 1e9, 3), "best": round(total_bytes / best / 1e9, 3), "ms_per_step": round(mean * 1e3, 3), "iters": iters}
        except Exception as e:
            res[path] = {"unavailable": str(e)}
    plain = res.get("plain", {})
    return {"value": plain.get("value"), "unit": "GB/s", "cores": threads, "kind": "reference", "path": "plain",
            "sample": f"whole step ({len(mats)} mul_mats, N={N}) x {plain.get('iters')} iterations on ggml_backend_cpu, AVX2 vec_dot, plain CPU "
                      f"buffer (no repack), ggml threadpool (no OpenMP); mean step {plain.get('ms_per_step')} ms",
            "steps_per_s": round(1e3 / plain["ms_per_step"], 2) if plain.get("ms_per_step") else None,
            "repack": dict(res.get("repack", {}), path="repack", note="weights in the CPU_REPACK buffer type (block_q4_Kx8 + ggml_gemv/gemm_q4_K_8x8_q8_K, "
                                                                         "L/ggml/src/ggml-cpu/repack.cpp:1844-1849): what llama-bench runs by default on this host")}


def run_reference(args, rank):
    if rank != 0:
        return
    from oracle import ref
    mats = workload_mats(args.workload)
    N = args.batch
    total = sum(algo_bytes(M, K, N) for _, M, K in mats)
    threads = host_threads()
    _, mean, _ = ref.mul_mat_set_ex([(M, K, 0) for _, M, K in mats], N, threads, max(args.warmup, 1), max(args.steps, 1), False)
    val = total / mean / 1e9
    line = {"impl": "reference", "metric": "q4k_gemv_hbm_gbs", "value": round(val, 3), "unit": "GB/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(mean * 1e3, 4),
            "higher_is_better": True, "scaling": "strong" if args.workload == "ffn70b" else "weak",
            "vs_baseline": None, "dtype": "int8 x uint4 -> int32, f32 scales", "data": "synthetic (random valid block_q4_K bytes)",
            "config": config_obj(args.workload, N, len(mats), max(args.gpus, 1)),
            "cpu_baseline": {"value": round(val, 3), "unit": "GB/s", "cores": threads, "kind": "reference", "path": "plain",
                             "sample": f"whole step x {args.steps} iterations; ggml_backend_cpu / ggml_compute_forward_mul_mat / AVX2 vec_dot, plain CPU buffer"},
            "e2e": {"value": round(val, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "decode_tok_s": round(1.0 / mean, 2) if args.workload == "tinyllama-token" else None,
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ per-shape tables (N = 1 GPU)
def gemv_shape_table(torch, q4k, device, shapes, peak, steps=20, cpu=True):
    """Per-shape decode figures (quantize + GEMV through q4k_mul_mat), rotating through > 2 x L2 of distinct weight copies so the
    bytes come from HBM; each with its fraction of the measured and the nominal HBM peak and the reference's CPU figures."""
    out = []
    for M, K, N in shapes:
        wbytes = M * (K // 256) * 144
        copies = max(2, (2 * L2_BYTES + wbytes) // wbytes + 1)
        ms = MatSet(torch, q4k, [(f"copy{i}", M, K) for i in range(copies)], N, device, seed=99)
        reps = max(1, -(-16 // copies))                    # >= 16 back-to-back launches per graph replay: the replay gap is not the kernel's

        def enqueue():
            for _ in range(reps):
                ms.enqueue_device()
        g = graph_of(torch, enqueue)
        for _ in range(3):
            g.replay()
        t = timed_steps(torch, g.replay, steps, 2, lambda: None) / (steps * copies * reps)
        b = algo_bytes(M, K, N)
        gbs = b / t / 1e6
        row = {"M": M, "K": K, "N": N, "us": round(t * 1e3, 3), "gbs": round(gbs, 1), "weight_copies": copies, "bound": "hbm", **hbm_fracs(gbs, peak)}
        if cpu:
            row["cpu"] = cpu_figures_for_shape(M, K, N, "GB/s")
        out.append(row)
        del ms, g
        torch.cuda.empty_cache()
    return out


def gemm_table(torch, q4k, device, peak_hbm, i8, cpu=True):
    """Prefill figures: quantizer + tcgen05 kind::i8 GEMM (or the batched decode kernel for mid N) through q4k_mul_mat.  int8 ops
    = 2*M*N*K (op_flops of L/tests/test-backend-ops.cpp:3308-3311; the exact hi/lo split's extra MMA work is NOT counted).
    Fractions: of the measured kind::i8 MMA rate (N = 256 instruction shape), of the exactness ceiling (half of it) and of nominal 4500."""
    out = {"shapes": [], "int8_peak": i8}
    peak_i8 = i8["ts_n256"]

    def run(mats, N, iters):
        ms = MatSet(torch, q4k, mats, N, device, seed=5)
        for _ in range(2):
            ms.enqueue_device()
        t = timed_steps(torch, ms.enqueue_device, iters, 1, lambda: None) / iters
        ops = sum(2.0 * M * K * N for _, M, K in mats)
        del ms
        torch.cuda.empty_cache()
        return t, ops

    def fr(tops):
        return {"frac_measured_i8": round(tops / peak_i8, 4), "frac_exact_ceiling": round(tops / (peak_i8 / 2), 4), "frac_nominal_4500": round(tops / 4500.0, 4)}

    for M, K, N in [(4096, 14336, 512), (14336, 4096, 512), (5632, 2048, 512), (4096, 14336, 32), (14336, 4096, 32)]:
        t, ops = run([("m", M, K)], N, 10)
        tops = ops / t / 1e9
        row = {"M": M, "K": K, "N": N, "us": round(t * 1e3, 1), "tops": round(tops, 1)}
        if N <= 64:       # HBM-bound by roofline: reported against HBM
            gbs = algo_bytes(M, K, N) / t / 1e6
            row.update({"bound": "hbm", "gbs": round(gbs, 1), **hbm_fracs(gbs, peak_hbm)})
        else:
            row.update({"bound": "tensor", **fr(tops)})
        if cpu:
            row["cpu"] = cpu_figures_for_shape(M, K, N, "GB/s" if N <= 64 else "TOPS")
        out["shapes"].append(row)
    t, ops = run(tinyllama_token_matmuls(), 512, 3)
    tops = ops / t / 1e9
    out["tinyllama_pp512"] = {"ms_per_pass": round(t, 3), "tops": round(tops, 1), "tok_s": round(512 / t * 1e3, 0), "int8_ops_per_pass": ops, **fr(tops)}
    return out


def q6k_shape_table(torch, q4k, device, shapes, peak, steps=20):
    """Per-shape Q6_K decode figures (q6k_mul_mat: quantizer fused), weight copies rotated through > 2 x L2."""
    out = []
    gen = torch.Generator(device=device)
    gen.manual_seed(5)
    for M, K, N in shapes:
        wbytes = M * (K // 256) * 210
        copies = max(2, (2 * L2_BYTES + wbytes) // wbytes + 1)
        ws = [rand_q6k(torch, M, K, device, gen) for _ in range(copies)]
        x = torch.randn((N, K), device=device, generator=gen)
        dst = torch.empty((N, M), device=device)
        wsb = torch.empty(max(16, N * (K // 256) * 292), dtype=torch.uint8, device=device)
        reps = max(1, -(-16 // copies))

        def enqueue():
            for _ in range(reps):
                for w in ws:
                    q4k.mul_mat_q6k(w, x, dst, wsb)
        g = graph_of(torch, enqueue)
        t = timed_steps(torch, g.replay, steps, 2, lambda: None) / (steps * copies * reps)
        gbs = (wbytes + N * K * 4 + M * N * 4) / t / 1e6
        out.append({"M": M, "K": K, "N": N, "us": round(t * 1e3, 3), "gbs": round(gbs, 1), "weight_copies": copies, "bound": "hbm", **hbm_fracs(gbs, peak)})
        del ws, g
        torch.cuda.empty_cache()
    return out


def q6k_prefill_obj(torch, q4k, device, peaks_json):
    """Q6_K prefill (q6k_mul_mat at N = 512: quantizer + tcgen05 kind::f16 GEMM, csrc/gemm_q6k_tc.cu) per shape, and the whole
    TinyLlama-1.1B Q4_K_M prompt pass — all 155 weight mul_mats, Q4_K and Q6_K — as one q4k_plan at N = 512.
    ops = 2*M*N*K.  The exact integers take two fp16 MMAs per weight: ceiling = half the dense fp16 rate."""
    gen = torch.Generator(device=device)
    gen.manual_seed(9)
    bf16 = float(peaks_json.get("bf16_tflops", 0.0)) or None
    out = {"shapes": [], "fp16_dense_tflops": bf16, "fp16_source": "MEASURED_PEAKS.json bf16_tflops (cuBLAS burst)" if bf16 else "unavailable"}
    L = q4k.load()
    N = 512
    for M, K in [(32000, 2048), (2048, 5632), (256, 2048), (4096, 14336)]:
        w = rand_q6k(torch, M, K, device, gen)
        x = torch.randn((N, K), device=device, generator=gen)
        dst = torch.empty((N, M), device=device)
        ws = torch.empty(L.q6k_mul_mat_workspace_bytes(K, N), dtype=torch.uint8, device=device)

        def call():
            q4k.mul_mat_q6k(w, x, dst, ws)
        for _ in range(2):
            call()
        t = timed_steps(torch, call, 10, 1, lambda: None) / 10
        tops = 2.0 * M * K * N / t / 1e9
        row = {"M": M, "K": K, "N": N, "us": round(t * 1e3, 1), "tops": round(tops, 1), "bound": "tensor"}
        if bf16:
            row["frac_exact_ceiling"] = round(tops / (bf16 / 2), 4)
        out["shapes"].append(row)
        del w, x, dst, ws
        torch.cuda.empty_cache()
    mats = tinyllama_full_token_matmuls()
    ids = shared_input_ids([(n, M, K) for n, M, K, _ in mats])
    plan = q4k.GraphPlan(torch.device(device).index or 0)
    keep = []
    for (name, M, K, q6), iid in zip(mats, ids):
        w = (rand_q6k if q6 else rand_q4k)(torch, M, K, device, gen)
        keep.append(w)
        plan.add(w, N, iid, q6k=q6)
    plan.finalize()
    for iid in set(ids):
        plan.input(iid).normal_(generator=gen)
    t = timed_steps(torch, plan.run, 3, 2, lambda: None) / 3
    ops = sum(2.0 * M * K * N for _, M, K, _ in mats)
    out["tinyllama_pp512_full"] = {"workload": "every weight mul_mat of a TinyLlama-1.1B Q4_K_M prompt pass at 512 tokens: 134 Q4_K + 21 Q6_K",
                                   "mul_mats": len(mats), "launches_per_pass": int(plan.launches), "ms_per_pass": round(t, 3),
                                   "tops": round(ops / t / 1e9, 1), "tok_s": round(512 / t * 1e3, 0)}
    del plan, keep
    torch.cuda.empty_cache()
    return out


def plugin_e2e(mats_typed, ids, iters):
    """`e2e_plugin`: the same mul_mat set as ONE ggml graph on the backend plugin, through the reference's public backend API
    (oracle/_ref's libggml drives it: tensor_set_async -> graph_compute_async | graph_plan_compute -> tensor_get_async ->
    synchronize), wall-clock per step.  The reference's libraries are the HOST here, not the thing measured."""
    from oracle import ref
    plugin = os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libggml-q4k-b200.so")
    if not (ref.available() and os.path.exists(plugin)):
        return {"unavailable": "oracle/_ref or the plugin is not built"}
    out = {}
    for key, use_plan in (("graph_compute", False), ("graph_plan_compute", True)):
        r = ref.plugin_graph_set(plugin, [(M, K, q6) for _, M, K, q6 in mats_typed], ids, 1, use_plan, 5, iters)
        out[key] = {"ms_per_step": round(r["mean_s"] * 1e3, 5), "best_ms": round(r["best_s"] * 1e3, 5)}
        out["h2d_bytes_per_step"], out["d2h_bytes_per_step"], out["pinned_host_buffer"] = r["h2d_bytes"], r["d2h_bytes"], r["pinned"]
    return out


def decode_token_obj(torch, q4k, device, steps, warmup, peak, peak_src, no_cpu):
    """configs[1] at N = 1: the 134 Q4_K mul_mats of a TinyLlama-1.1B Q4_K_M decode token as ONE q4k_plan, with its own roofline,
    e2e (q4k_plan_run_host), e2e_plugin (the same graph through the ggml backend API on the plugin) and cpu_baseline."""
    mats = tinyllama_token_matmuls()
    ms = MatSet(torch, q4k, mats, 1, device, seed=1234)
    plan = ms.make_plan()
    t = timed_steps(torch, plan.run, steps, warmup, lambda: None) / steps

    def step_host():
        plan.run_host(ms.plan_in_host, ms.plan_out_host)
        torch.cuda.current_stream().synchronize()
    th = timed_steps(torch, step_host, steps, warmup, lambda: None) / steps
    total = ms.algo_bytes
    launches = int(plan.launches)
    per_launch_us = t * 1e3 / launches
    achieved = total / launches / per_launch_us / 1e3
    tr = NCU_TRAFFIC["tinyllama-token"]
    obj = {"workload": workload_name("tinyllama-token"), "mul_mats": len(mats), "launches_per_step": launches,
           "algorithmic_bytes": total, "weight_bytes": ms.weight_bytes,
           "value": round(total / t / 1e6, 2), "unit": "GB/s", "ms_per_step": round(t, 5), "decode_tok_s": round(1e3 / t, 1),
           **hbm_fracs(total / t / 1e6, peak),
           "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                        "traffic": tr["bytes"], "traffic_kernel": tr["kernel"], "traffic_source": tr["source"], "peak_source": peak_src,
                        "kernel": "gemv_q4k_mma_kernel (Q8_K quantizer fused into the launch)",
                        "bytes_per_launch": round(total / launches), "us_per_launch": round(per_launch_us, 3),
                        "note": "launch / latency-bound: 0.05-1.0 us of HBM time per launch against a ~2 us dependent-launch floor"},
           "e2e": {"value": round(total / th / 1e6, 2), "unit": "GB/s", "ms_per_step": round(th, 5), "decode_tok_s": round(1e3 / th, 1),
                   "h2d_bytes_per_step": plan.input_bytes, "d2h_bytes_per_step": plan.output_bytes,
                   "path": "q4k_plan_run_host: packed pinned inputs up, four graphs, packed results down; copies pipelined under the kernels"}}
    del plan, ms
    torch.cuda.empty_cache()
    try:
        ids = shared_input_ids(mats)
        obj["e2e_plugin"] = plugin_e2e([(n, M, K, False) for n, M, K in mats], ids, max(steps, 20))
    except Exception as e:
        obj["e2e_plugin"] = {"unavailable": f"{type(e).__name__}: {e}"}
    if not no_cpu:
        try:
            obj["cpu_baseline"] = cpu_baseline_obj(mats, 1, total, budget_s=8.0)
        except Exception as e:
            obj["cpu_baseline"] = {"value": None, "unit": "GB/s", "cores": 0, "kind": "reference", "sample": f"unavailable: {e}"}
    return obj


def full_token_extra(torch, q4k, device, steps, no_cpu, peak):
    """the whole Q4_K_M token (Q4_K + Q6_K mul_mats, 155 of them) as ONE q4k_plan on the GPU, through the
    plugin, and on the reference's CPU backend (plain and repack)."""
    mats = tinyllama_full_token_matmuls()
    gen = torch.Generator(device=device)
    gen.manual_seed(77)
    ids = shared_input_ids([(n, M, K) for n, M, K, _ in mats])
    plan = q4k.GraphPlan(torch.device(device).index or 0)
    keep, nbytes = [], 0
    for (name, M, K, q6), iid in zip(mats, ids):
        w = (rand_q6k if q6 else rand_q4k)(torch, M, K, device, gen)
        keep.append(w)
        plan.add(w, 1, iid, q6k=q6)
        nbytes += w.numel() + K * 4 + M * 4
    plan.finalize()
    for iid in set(ids):
        plan.input(iid).normal_(generator=gen)
    ms = timed_steps(torch, plan.run, steps, 5, lambda: None) / steps
    in_host = torch.zeros(plan.input_bytes, dtype=torch.uint8).pin_memory()
    out_host = torch.zeros(plan.output_bytes, dtype=torch.uint8).pin_memory()
    in_host.copy_(plan._in)

    def step_host():
        plan.run_host(in_host, out_host)
        torch.cuda.current_stream().synchronize()
    ms_host = timed_steps(torch, step_host, steps, 5, lambda: None) / steps
    out = {"workload": "tinyllama-1.1b-q4_k_m decode token, every weight mul_mat: 134 Q4_K + 21 Q6_K (attn_v / ffn_down of 10 layers, output.weight)",
           "mul_mats": len(mats), "launches_per_step": int(plan.launches), "algorithmic_bytes": nbytes,
           "ms_per_step": round(ms, 5), "gbs": round(nbytes / ms / 1e6, 1), "decode_tok_s": round(1e3 / ms, 1), **hbm_fracs(nbytes / ms / 1e6, peak),
           "e2e": {"ms_per_step": round(ms_host, 5), "gbs": round(nbytes / ms_host / 1e6, 1), "decode_tok_s": round(1e3 / ms_host, 1),
                   "h2d_bytes_per_step": plan.input_bytes, "d2h_bytes_per_step": plan.output_bytes}}
    del plan, keep
    torch.cuda.empty_cache()
    try:
        out["e2e_plugin"] = plugin_e2e(mats, ids, max(steps, 20))
    except Exception as e:
        out["e2e_plugin"] = {"unavailable": f"{type(e).__name__}: {e}"}
    if not no_cpu:
        try:
            out["cpu_reference"] = cpu_baseline_obj([(n, M, K) for n, M, K, _ in mats], 1, nbytes, budget_s=5.0, q6=[q for *_, q in mats])
        except Exception as e:
            out["cpu_reference"] = {"unavailable": str(e)}
    return out


# ------------------------------------------------------------------------------------------------ main
def llama_bench_obj(reps=2):
    """The reference's own headline tool (README.md:189-193: llama-bench -p 512 -n 128) on a random-init TinyLlama-1.1B Q4_K_M
    GGUF (tools/make_gguf.py): on the host cores alone, and with the plugin loaded through GGML_BACKEND_PATH (-ngl 99; the KV cache
    and the attention stay on the host: -nkvo 1).  Scheduler-bound end-to-end figures of the whole llama graph, not kernel numbers."""
    import subprocess
    import tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "llama-bench")
    plugin = os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libggml-q4k-b200.so")
    if not (os.path.exists(exe) and os.path.exists(plugin)):
        return {"unavailable": "oracle/_ref/llama-bench or the plugin is not built"}
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import make_gguf
        gguf = os.path.join(tempfile.gettempdir(), "q4k_b200_tinyllama_random_q4_k_m.gguf")
        info = make_gguf.write(gguf)
        nt = min(host_threads(), 16)
        out = {"model": "tinyllama-1.1b-shaped, random-init, Q4_K_M tensor types, no_vocab", "gguf_bytes": info["file_bytes"], "threads": nt,
               "command": "llama-bench -p 512 -n 128", "unit": "tokens/s"}

        def run(extra, env):
            r = subprocess.run([exe, "-m", gguf, "-p", "512", "-n", "128", "-t", str(nt), "-r", str(reps), "-o", "json"] + extra,
                               capture_output=True, text=True, timeout=300, env=env, cwd=os.path.dirname(exe))
            try:
                rows = json.loads(r.stdout[r.stdout.index("["):])
                return {("pp512" if row["n_prompt"] else "tg128"): round(row["avg_ts"], 2) for row in rows}
            except Exception:
                return {"unavailable": (r.stderr or r.stdout).strip().splitlines()[-2:]}

        base_env = {k: v for k, v in os.environ.items() if k not in ("GGML_BACKEND_PATH", "Q4K_B200_CHEAP_OPS")}
        out["cpu_reference"] = run([], base_env)
        out["plugin"] = dict(run(["-ngl", "99", "-nkvo", "1", "-sm", "none"], dict(base_env, GGML_BACKEND_PATH=plugin)),
                             flags="-ngl 99 -nkvo 1 -sm none", graph_splits_per_token=46)
        out["plugin_mul_mat_only"] = dict(run(["-ngl", "99", "-nkvo", "1", "-sm", "none", "-p", "0"],
                                              dict(base_env, GGML_BACKEND_PATH=plugin, Q4K_B200_CHEAP_OPS="0")),
                                          flags="Q4K_B200_CHEAP_OPS=0", graph_splits_per_token=266)
        os.remove(gguf)
        return out
    except Exception as e:                   # a reported extra, never a dependency of the GPU path
        return {"unavailable": f"{type(e).__name__}: {e}"[:300]}


def run_ours(args, rank, world, local_rank):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the Q4_K path has no CPU fallback")
    import q4k_b200
    lib = q4k_b200.load()
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if not lib.q4k_device_ok(local_rank):
        raise SystemExit("bench.py: device is not compute capability 10.x (sm_100a build)")
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if dist is not None:
            dist.barrier()

    N = args.batch
    sharded = args.workload == "ffn70b"
    mats = workload_mats(args.workload)
    from q4k_b200.shard import shard_rows
    row_range = (lambda M: shard_rows(M, world)[rank]) if (sharded and world > 1) else None
    ms = MatSet(torch, q4k_b200, mats, N, device, seed=1234 + (rank if row_range else 0), row_range=row_range)
    total_bytes = sum(algo_bytes(M, K, N) for _, M, K in mats) * (1 if sharded else world)

    shards, gather_kind, plan = None, None, None
    if sharded and world > 1:
        from q4k_b200.shard import RowShardedQ4K
        gen = torch.Generator(device=device)
        gen.manual_seed(777)                                   # activations are replicated: same numbers on every rank
        shards = []
        for it in ms.items:
            it["x"] = torch.randn((N, it["K"]), device=device, generator=gen)
            shards.append(RowShardedQ4K(it["w"], it["M_full"], it["K"], gather=args.gather))
        gather_kind = shards[0].gather + (f" [{shards[0].peer_note}]" if shards[0].peer_note else "")

        def enqueue():
            for it, sh in zip(ms.items, shards):
                it["full"] = sh.mul_mat(it["x"])
        # parity BEFORE timing: every rank's rows of every matrix, sampled, against the oracle (rank 0 checks what IT holds)
        results = []
        for it, sh in zip(ms.items, shards):
            results.append(sh.mul_mat(it["x"]).clone())
        torch.cuda.synchronize()
        barrier()
        parity_ok, parity = parity_before_timing(torch, ms.items, results, N, world, rank, dist)
        del results
        step, one_step, steps_per_call = enqueue, enqueue, 1
        n_cap = q4k_b200.load().q4k_kernel_launches()
        enqueue()
        launches_eager_step = q4k_b200.load().q4k_kernel_launches() - n_cap     # a graph replay launches the same kernels without passing the library's counter
        torch.cuda.synchronize()
        barrier()
        if args.graph and shards[0].gather in ("ll", "peer"):        # NCCL collectives are left out of graph capture
            try:                                               # the whole step (GEMVs + cross-GPU exchange) as one CUDA graph
                if shards[0].needs_even_capture:               # two steps per graph: the host-chosen result buffers keep alternating
                    step, steps_per_call = graph_of(torch, lambda: (enqueue(), enqueue())).replay, 2
                else:
                    step = graph_of(torch, enqueue).replay
                gather_kind += f", CUDA graph of {steps_per_call} step(s)"
            except Exception as e:
                torch.cuda.synchronize()
                gather_kind += f", stream launches (graph capture failed: {type(e).__name__})"
    else:
        steps_per_call = 1
        if args.plan:
            plan = ms.make_plan()                                  # the library's own graph plan (q4k_plan_*)
            step = plan.run
            step()
            torch.cuda.synchronize()
            results = [plan.output(it["plan_index"]) for it in ms.items]
        else:
            g = graph_of(torch, ms.enqueue_device)
            step = g.replay
            step()
            torch.cuda.synchronize()
            results = [it["dst"] for it in ms.items]
        # the tinyllama token has 134 matrices: check a spread of 24 of them
        sel = list(range(len(ms.items))) if len(ms.items) <= 24 else sorted(set(range(0, len(ms.items), max(1, len(ms.items) // 24))))
        parity_ok, parity = parity_before_timing(torch, [ms.items[i] for i in sel], [results[i] for i in sel], N, 1, 0, None)
        one_step = step
    n0 = lib.q4k_kernel_launches()
    step()
    launches_per_step = (lib.q4k_kernel_launches() - n0) // steps_per_call
    if launches_per_step == 0 and sharded and world > 1:     # graph replay of the sharded step: the kernels counted when the same step ran eagerly
        launches_per_step = launches_eager_step
    torch.cuda.synchronize()

    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    steps_timed = args.steps

    def k_steps(k):                                            # exactly k steps: graph replays of steps_per_call steps + single steps
        for _ in range(k // steps_per_call):
            step()
        for _ in range(k % steps_per_call):
            one_step()
    ms_total = timed_region(torch, lambda: k_steps(args.steps), lambda: k_steps(args.warmup + args.warmup % steps_per_call), barrier)
    t0 = torch.tensor([ms_total], device=device, dtype=torch.float64)
    if dist is not None:                                       # every rank must replay the SAME number of steps below
        dist.all_reduce(t0, op=dist.ReduceOp.MAX)
    per_step_ms0 = float(t0[0]) / steps_timed
    sus_n, sus_ms = sustained(torch, step, 0.5, per_step_ms0 * steps_per_call, barrier)
    # end to end: pinned host activations in, host results out, every step
    if sharded and world > 1:
        for it in ms.items:
            it["x_host"] = it["x"].cpu().pin_memory()
            it["x_stage"] = torch.empty_like(it["x"])
            it["full_host"] = torch.zeros((N, it["M_full"]), dtype=torch.float32).pin_memory()
        ms.h2d = sum(it["x_host"].numel() * 4 for it in ms.items)
        ms.d2h = sum(it["full_host"].numel() * 4 for it in ms.items)

        def step_host():
            for it, sh in zip(ms.items, shards):
                it["x_stage"].copy_(it["x_host"], non_blocking=True)
                full = sh.mul_mat(it["x_stage"])
                it["full_host"].copy_(full, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        e2e_path = "per mul_mat: pinned x up, sharded mul_mat + exchange, full result down; one stream sync per step"
    elif plan is not None:
        def step_host():                                       # packed H2D, the graph(s), packed D2H (pipelined inside q4k_plan_run_host)
            plan.run_host(ms.plan_in_host, ms.plan_out_host)
            torch.cuda.current_stream().synchronize()
        e2e_path = "q4k_plan_run_host: packed pinned inputs up, the plan's graphs, packed results down; copies pipelined under the kernels on a side stream"
    else:
        for it in ms.items:
            it["x_host"] = it["x"].cpu().pin_memory()
            it["dst_host"] = torch.zeros((N, it["M"]), dtype=torch.float32).pin_memory()
            it["x_stage"] = torch.empty_like(it["x"])
        ms.h2d = sum(it["x_host"].numel() * 4 for it in ms.items)
        ms.d2h = sum(it["dst_host"].numel() * 4 for it in ms.items)

        def step_host():
            st = torch.cuda.current_stream().cuda_stream
            for it in ms.items:
                rc = lib.q4k_mul_mat_host_async(it["w"].data_ptr(), it["M"], it["K"], it["w"].stride(0), it["x_host"].data_ptr(), N,
                                                it["dst_host"].data_ptr(), it["x_stage"].data_ptr(), it["dst"].data_ptr(), ms.ws.data_ptr(), ms.ws.numel(), st)
                if rc != 0:
                    raise RuntimeError(f"q4k_mul_mat_host_async rc={rc}")
            torch.cuda.current_stream().synchronize()
        e2e_path = "q4k_mul_mat_host_async per mul_mat"
    ms_e2e = timed_steps(torch, step_host, args.steps, args.warmup, barrier)
    clk = clocks.stop() if rank == 0 else None

    t = torch.tensor([ms_total, ms_e2e, sus_ms], device=device, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, sus_ms = float(t[0]), float(t[1]), float(t[2])
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peak, peak_src, peaks_json = measured_peaks()
    per_step_ms = ms_total / steps_timed
    value = total_bytes / (per_step_ms * 1e-3) / 1e9
    e2e_val = total_bytes / (ms_e2e / args.steps * 1e-3) / 1e9
    sus_val = total_bytes / (sus_ms / (sus_n * steps_per_call) * 1e-3) / 1e9
    # dominant kernel = the decode GEMV of a mul_mat: algorithmic bytes per mul_mat / its average slot inside the timed region (per rank)
    n_slots = len(ms.items) if sharded else int(launches_per_step)
    per_slot_bytes = ms.algo_bytes / n_slots
    per_slot_us = per_step_ms * 1e3 / n_slots
    achieved = per_slot_bytes / per_slot_us / 1e3
    tr = NCU_TRAFFIC.get(args.workload) if world == 1 else None
    if args.workload == "ffn70b":
        kernel = ("gemv_q4k_tma_kernel<1, 2> for 28672x8192 (large-matrix decode launch, quantizer fused) and gemv_q4k_n1s_kernel for 8192x28672 "
                  "(activations quantized once by quantize_q8_K_n1s_kernel and streamed next to the weights)" if world == 1
                  else "gemv_q4k_mma_kernel / gemv_q4k_tma_kernel on this rank's row shard + the cross-GPU exchange")
    else:
        kernel = "gemv_q4k_mma_kernel (Q8_K quantizer fused into the launch)"
    line = {"metric": "q4k_gemv_hbm_gbs", "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": steps_timed,
            "warmup": args.warmup, "ms_per_step": round(per_step_ms, 5), "higher_is_better": True,
            "scaling": "strong" if sharded else "weak", "vs_baseline": None,
            "dtype": "int8 x uint4 -> int32, f32 scales", "data": "synthetic (random valid block_q4_K bytes, N(0,1) activations)",
            "config": config_obj(args.workload, N, len(ms.items), world),
            "run_info": {"weight_bytes_per_step_per_gpu": ms.weight_bytes,
                         "l2": f"no flush needed: each step streams {ms.weight_bytes >> 20} MiB of distinct weights per GPU (> 126 MiB L2)"
                               if ms.weight_bytes > L2_BYTES else "WARNING: per-GPU weights fit in L2",
                         "launch": (f"q4k_plan_run: one CUDA graph of {int(launches_per_step)} launches" if plan is not None
                                    else "CUDA graph replay of the C-ABI launches") if not (sharded and world > 1) else f"gather={gather_kind}",
                         "e2e_path": e2e_path, "launches_per_step": int(launches_per_step)},
            "decode_tok_s": round(1e3 / per_step_ms, 1) if not sharded else None,
            "pct_of_nominal_8TBs": round(100 * value / world / HBM_NOMINAL_GBS, 2),
            "parity_checked": bool(parity_ok), "parity": parity,
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "frac_nominal_8TBs": round(achieved / HBM_NOMINAL_GBS, 4),
                         "traffic": tr["bytes"] if tr else None, "traffic_kernel": tr["kernel"] if tr else None, "traffic_source": tr["source"] if tr else None,
                         "peak_source": peak_src, "kernel": kernel,
                         "bytes_per_launch": round(per_slot_bytes), "us_per_launch": round(per_slot_us, 3),
                         "definition": "algorithmic bytes of one mul_mat on this rank / (device time of the timed region / mul_mats in it)"},
            "sustained": {"value": round(sus_val, 2), "unit": "GB/s", "steps": sus_n * steps_per_call, "seconds": round(sus_ms / 1e3, 3),
                          "note": "the same step repeated for ~0.5 s, device-timed: the K-step figure is not a burst"},
            "e2e": {"value": round(e2e_val, 2), "unit": "GB/s", "h2d_bytes_per_step": ms.h2d, "d2h_bytes_per_step": ms.d2h,
                    "ms_per_step": round(ms_e2e / args.steps, 5),
                    "decode_tok_s": round(1e3 / (ms_e2e / args.steps), 1) if not sharded else None},
            "gpu_launches": int(launches_per_step * steps_timed), "clocks": clk}
    if world == 1 and not args.no_cpu:
        try:
            line["cpu_baseline"] = cpu_baseline_obj(mats, N, total_bytes, budget_s=8.0)
        except Exception as e:       # the baseline is a reported number, never a dependency of the GPU path
            line["cpu_baseline"] = {"value": None, "unit": "GB/s", "cores": 0, "kind": "reference", "sample": f"unavailable: {e}"}
    if world == 1 and args.shapes:
        del ms, plan
        torch.cuda.empty_cache()
        cpu = not args.no_cpu
        if args.workload == "ffn70b":
            line["decode_token"] = decode_token_obj(torch, q4k_b200, device, max(args.steps, 20), args.warmup, peak, peak_src, args.no_cpu)
        line["shapes"] = gemv_shape_table(torch, q4k_b200, device,
                                          [(5632, 2048, 1), (4096, 14336, 1), (14336, 4096, 1), (4096, 14336, 8), (14336, 4096, 8),
                                           (4096, 14336, 16), (28672, 8192, 1), (8192, 28672, 1), (28672, 8192, 8), (8192, 28672, 8), (28672, 8192, 16)], peak, cpu=cpu)
        line["prefill"] = gemm_table(torch, q4k_b200, device, peak, int8_peaks((clk or {}).get("sm_mhz")), cpu=cpu)
        line["full_q4_k_m_token"] = full_token_extra(torch, q4k_b200, device, max(args.steps, 20), args.no_cpu, peak)
        if cpu:
            line["llama_bench"] = llama_bench_obj()
        try:
            line["q6k_prefill"] = q6k_prefill_obj(torch, q4k_b200, device, peaks_json)
        except Exception as e:
            line["q6k_prefill"] = {"unavailable": f"{type(e).__name__}: {e}"}
        line["q6k_shapes"] = q6k_shape_table(torch, q4k_b200, device, [(2048, 5632, 1), (32000, 2048, 1), (14336, 4096, 1), (28672, 8192, 1)], peak)
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ffn70b", choices=["tinyllama-token", "ffn70b"])
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--no-graph", dest="graph", action="store_false", help="sharded path: plain stream launches")
    ap.add_argument("--gather", default="auto", choices=["auto", "nccl", "ll", "peer"], help="row-shard output exchange (N > 1 GPUs)")
    ap.add_argument("--no-plan", dest="plan", action="store_false", help="single GPU: one q4k_mul_mat per launch instead of the graph plan")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU legs")
    ap.add_argument("--shapes", action="store_true", default=True, help="add the decode_token object, the per-shape tables and the prefill figures")
    ap.add_argument("--no-shapes", dest="shapes", action="store_false")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, world, local_rank)
        if world > 1:                       # symmetric-memory / NCCL teardown order at interpreter exit is not ours to debug
            sys.stdout.flush()
            os._exit(0)


if __name__ == "__main__":
    main()
