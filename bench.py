#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the Q4_K x Q8_K mul_mat hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one pass of the hot path over one batch of synthetic input:

  N = 1 (default)  workload "tinyllama-1.1b-q4_k_m-decode-token" (BASELINE.json configs[1]): every
                   GGML_TYPE_Q4_K mul_mat of ONE decode token of TinyLlama-1.1B Q4_K_M, batch 1, in llama
                   graph order — 134 mul_mats, 477 167 616 B of Q4_K weights (the Q6_K tensors of the
                   Q4_K_M mix are not on this path: SURVEY.md §8d C2).  Each mul_mat = quantize_row_q8_K of
                   its F32 activations + the GEMV, as ggml_compute_forward_mul_mat does
                   (L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401).
  N > 1 (torchrun) workload "llama3-70b-ffn-rowshard" (configs[4]): the 28672x8192 and 8192x28672 Q4_K
                   matrices, batch 1, weight rows sharded over the N ranks, F32 output slices all-gathered
                   (SURVEY.md §8e).  Fixed total work -> "scaling": "strong".

`value` = algorithmic bytes of the step (M*K/256*144 + N*K*4 + M*N*4 per mul_mat) / device
time, inputs resident in HBM.  `e2e` = the same through the host-buffer C-ABI entry (q4k_mul_mat_host_async):
pinned host activations in, host results out, copies inside the timed region.  `--impl reference` times the
UNMODIFIED reference ggml CPU code (oracle/_ref, built by oracle/Makefile) on the host cores on the same
workload.  oracle/ is used here only as that CPU baseline, never on the measured GPU path.
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
# dram__bytes_read.sum + dram__bytes_write.sum of one gemv_q4k_mma_kernel launch from the round's `ncu --set full` capture
NCU_TRAFFIC = {"bytes": 6562304, "note": "ncu --set full of the 5632x2048 N=1 launch (profiles/r1_ncu_gemv_mma_c1_5632x2048.txt): "
                                         "6562304 B DRAM read + 0 B write per launch vs 6518784 B algorithmic; the 28672x8192 launch reads "
                                         "132.22 MB for 132.19 MB algorithmic (profiles/r1_ncu_gemv_mma_c5_28672x8192.txt)"}
Q6K_LAYERS = {0, 1, 4, 7, 10, 13, 16, 19, 20, 21}   # use_more_bits(i, 22), L/src/llama-quant.cpp:185-187


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


def ffn70b_matmuls():
    return [("ffn_up_70b", 28672, 8192), ("ffn_down_70b", 8192, 28672)]


def algo_bytes(M, K, N):
    return M * (K // 256) * 144 + N * K * 4 + M * N * 4


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", p
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)", {}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.thr = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
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
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ GPU arm
def rand_q4k(torch, M, K, device, gen):
    """Random VALID block_q4_K bytes (any byte pattern with finite fp16 d/dmin is a valid block):
    throughput-only weights, d = 2^-4, dmin = 2^-5.  Parity is tested elsewhere on really quantized weights."""
    nb = K // 256
    w = torch.randint(0, 256, (M, nb, 144), dtype=torch.uint8, device=device, generator=gen)
    w[..., 0] = 0x00; w[..., 1] = 0x2c; w[..., 2] = 0x00; w[..., 3] = 0x28
    return w.view(M, nb * 144)


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
            shared = N > 12 and self.items and self.items[-1]["iid"] == iid and self.items[-1]["K"] == K
            if shared:
                x = self.items[-1]["x"]
            dst = torch.zeros((N, r1 - r0), device=device)
            self.items.append(dict(name=name, M=r1 - r0, K=K, M_full=M, w=w, x=x, dst=dst, iid=iid, reuse=bool(shared)))
        kmax = max(K for _, _, K in mats)
        self.ws = torch.empty(max(q4k.workspace_bytes(kmax, N), 16), dtype=torch.uint8, device=device)
        self.algo_bytes = sum(algo_bytes(it["M"], it["K"], N) for it in self.items)
        self.weight_bytes = sum(it["w"].numel() for it in self.items)

    def enqueue_device(self):
        """One step with inputs resident in HBM: quantize + GEMV per mul_mat on the current stream."""
        st = self.torch.cuda.current_stream().cuda_stream
        ws, wsn, N, f = self.ws.data_ptr(), self.ws.numel(), self.N, self.lib.q4k_mul_mat
        for it in self.items:
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

    def make_host(self):
        t = self.torch
        for it in self.items:
            it["x_host"] = it["x"].cpu().pin_memory()
            it["dst_host"] = t.zeros((self.N, it["M"]), dtype=t.float32).pin_memory()
            it["x_stage"] = t.empty_like(it["x"])
        self.h2d = sum(it["x_host"].numel() * 4 for it in self.items)
        self.d2h = sum(it["dst_host"].numel() * 4 for it in self.items)

    def enqueue_host(self):
        """One step through the host-buffer C-ABI entry: H2D x, quantize, GEMV, D2H dst per mul_mat."""
        st = self.torch.cuda.current_stream().cuda_stream
        ws, wsn, N, f = self.ws.data_ptr(), self.ws.numel(), self.N, self.lib.q4k_mul_mat_host_async
        for it in self.items:
            rc = f(it["w"].data_ptr(), it["M"], it["K"], it["w"].stride(0), it["x_host"].data_ptr(), N,
                   it["dst_host"].data_ptr(), it["x_stage"].data_ptr(), it["dst"].data_ptr(), ws, wsn, st)
            if rc != 0:
                raise RuntimeError(f"q4k_mul_mat_host_async rc={rc} cuda={self.lib.q4k_last_cuda_error()}")


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


def graph_of(torch, enqueue):
    """Capture one step into a CUDA graph (the launches are ours; the graph removes host launch latency)."""
    enqueue()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        enqueue()
    return g


def gemm_table(torch, q4k, device, peaks):
    """Prefill (N = 512) figures: quantizer + tcgen05 kind::i8 GEMM through q4k_mul_mat.  int8 ops = 2*M*N*K
    (op_flops of L/tests/test-backend-ops.cpp:3308-3311; the hi/lo/min split's extra MMA work is NOT counted)."""
    out = {"shapes": []}
    int8_nominal = 4500.0                                   # dense int8 = dense fp8 rate on B200 (B200_PROFILING.md: 4.5 PFLOP/s)
    int8_from_bf16 = 2.0 * float(peaks.get("bf16_tflops", 0.0)) if peaks else 0.0

    def run(mats, N, iters):
        ms = MatSet(torch, q4k, mats, N, device, seed=5)
        for _ in range(2):
            ms.enqueue_device()
        t = timed_steps(torch, ms.enqueue_device, iters, 1, lambda: None) / iters
        ops = sum(2.0 * M * K * N for _, M, K in mats)
        del ms
        torch.cuda.empty_cache()
        return t, ops

    for M, K, N in [(4096, 14336, 512), (14336, 4096, 512), (4096, 14336, 32)]:
        t, ops = run([("m", M, K)], N, 10)
        out["shapes"].append({"M": M, "K": K, "N": N, "us": round(t * 1e3, 1), "tops": round(ops / t / 1e9, 1)})
    t, ops = run(tinyllama_token_matmuls(), 512, 3)
    tops = ops / t / 1e9
    out["tinyllama_pp512"] = {"ms_per_pass": round(t, 3), "tops": round(tops, 1), "tok_s": round(512 / t * 1e3, 0),
                              "int8_ops_per_pass": ops,
                              "frac_of_nominal_int8_4500": round(tops / int8_nominal, 4),
                              "frac_of_2x_measured_bf16": round(tops / int8_from_bf16, 4) if int8_from_bf16 else None}
    return out


def gemv_shape_table(torch, q4k, device, shapes, steps=20):
    """Per-shape GEMV figures (quantize + GEMV through q4k_mul_mat), rotating through > 2 x L2 of distinct
    weight copies so the bytes come from HBM.  Extra evidence; not the headline."""
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
        out.append({"M": M, "K": K, "N": N, "us": round(t * 1e3, 3), "gbs": round(b / t / 1e6, 1),
                    "weight_copies": copies})
        del ms, g
        torch.cuda.empty_cache()
    return out


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


def rand_q6k(torch, M, K, device, gen):
    """Random VALID block_q6_K bytes (any ql / qh / scales bytes are valid), d = 2^-8: throughput-only weights."""
    nb = K // 256
    w = torch.randint(0, 256, (M, nb, 210), dtype=torch.uint8, device=device, generator=gen)
    w[..., 208] = 0x00
    w[..., 209] = 0x1c
    return w.view(M, nb * 210)


def full_token_extra(torch, q4k, device, steps, no_cpu):
    """the whole Q4_K_M token (Q4_K + Q6_K mul_mats, 155 of them) as ONE q4k_plan on the GPU, and the
    same set on the reference's CPU backend.  Extra evidence next to the headline (which stays the 134 Q4_K mul_mats)."""
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
           "ms_per_step": round(ms, 5), "gbs": round(nbytes / ms / 1e6, 1), "decode_tok_s": round(1e3 / ms, 1),
           "e2e": {"ms_per_step": round(ms_host, 5), "gbs": round(nbytes / ms_host / 1e6, 1), "decode_tok_s": round(1e3 / ms_host, 1),
                   "h2d_bytes_per_step": plan.input_bytes, "d2h_bytes_per_step": plan.output_bytes}}
    if not no_cpu:
        try:
            from oracle import ref
            threads = host_threads()
            _, mean, _ = ref.mul_mat_set_typed([(M, K, q6) for _, M, K, q6 in mats], 1, threads, 1, 10)
            out["cpu_reference"] = {"ms_per_step": round(mean * 1e3, 3), "gbs": round(nbytes / mean / 1e9, 2), "decode_tok_s": round(1 / mean, 1),
                                    "cores": threads, "path": "ggml_backend_cpu: ggml_vec_dot_q4_K_q8_K / ggml_vec_dot_q6_K_q8_K (AVX2), 10 iterations"}
        except Exception as e:
            out["cpu_reference"] = {"unavailable": str(e)}
    del plan, keep
    torch.cuda.empty_cache()
    return out


def q6k_shape_table(torch, q4k, device, shapes, steps=20):
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
        out.append({"M": M, "K": K, "N": N, "us": round(t * 1e3, 3), "gbs": round((wbytes + N * K * 4 + M * N * 4) / t / 1e6, 1),
                    "weight_copies": copies})
        del ws, g
        torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------ reference arm
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def reference_step_seconds(mats, N, threads, warmup, iters):
    """The UNMODIFIED reference (oracle/_ref: ggml_mul_mat graph on ggml_backend_cpu -> ggml_compute_forward_mul_mat
    -> quantize_row_q8_K + ggml_vec_dot_q4_K_q8_K, AVX2 branch) on the whole mul_mat set.  mean seconds per step."""
    from oracle import ref
    if not ref.available():
        raise RuntimeError("oracle/_ref is not built (make -C oracle ref in the build container)")
    best, mean, _ = ref.mul_mat_set([(M, K) for _, M, K in mats], N, threads, warmup, iters)
    return best, mean


def cpu_baseline_obj(mats, N, total_bytes, budget_s=20.0):
    """Reference CPU code on this box's host cores, bounded to ~budget_s of CPU wall time."""
    threads = host_threads()
    t0 = time.time()
    best, mean = reference_step_seconds(mats, N, threads, 1, 2)
    iters = int(max(2, min(20, (budget_s - (time.time() - t0)) / max(mean, 1e-4))))
    best2, mean2 = reference_step_seconds(mats, N, threads, 1, iters)
    mean = min(mean, mean2)
    return {"value": round(total_bytes / mean / 1e9, 3), "unit": "GB/s", "cores": threads, "kind": "reference",
            "sample": f"whole step ({len(mats)} Q4_K mul_mats, N={N}) x {iters} iterations, ggml CPU backend, AVX2 vec_dot, "
                      f"plain CPU buffer (no repack), ggml threadpool (no OpenMP); mean step {mean * 1e3:.2f} ms, "
                      f"best {min(best, best2) * 1e3:.2f} ms",
            "steps_per_s": round(1.0 / mean, 2)}


def run_reference(args, rank):
    if rank != 0:
        return
    mats = tinyllama_token_matmuls() if args.workload == "tinyllama-token" else ffn70b_matmuls()
    N = args.batch
    total = sum(algo_bytes(M, K, N) for _, M, K in mats)
    threads = host_threads()
    _, mean = reference_step_seconds(mats, N, threads, max(args.warmup, 1), max(args.steps, 1))
    val = total / mean / 1e9
    line = {"impl": "reference", "metric": "q4k_gemv_hbm_gbs", "value": round(val, 3), "unit": "GB/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(mean * 1e3, 4),
            "higher_is_better": True, "scaling": "weak" if args.workload == "tinyllama-token" else "strong",
            "vs_baseline": None, "dtype": "int8 x uint4 -> int32, f32 scales", "data": "synthetic",
            "config": {"workload": workload_name(args.workload), "batch": N, "mul_mats": len(mats),
                       "threads": threads, "path": "ggml_backend_cpu / ggml_compute_forward_mul_mat / AVX2 vec_dot"},
            "cpu_baseline": {"value": round(val, 3), "unit": "GB/s", "cores": threads, "kind": "reference",
                             "sample": f"whole step x {args.steps} iterations"},
            "e2e": {"value": round(val, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "decode_tok_s": round(1.0 / mean, 2) if args.workload == "tinyllama-token" else None,
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_name(w):
    return {"tinyllama-token": "tinyllama-1.1b-q4_k_m-decode-token (configs[1]): 134 Q4_K mul_mats, batch 1",
            "ffn70b": "llama3-70b-ffn-rowshard (configs[4]): 28672x8192 + 8192x28672 Q4_K, batch 1"}[w]


# ------------------------------------------------------------------------------------------------ main
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
    mats = ffn70b_matmuls() if sharded else tinyllama_token_matmuls()
    if sharded:
        # distinct weight copies until every GPU streams > 2 x L2 per step: the bytes must come from HBM, not L2
        per_gpu = sum(M * (K // 256) * 144 for _, M, K in mats) / world
        copies = max(1, int(-(-2.2 * L2_BYTES // per_gpu)))
        mats = [(f"{n}.{c}", M, K) for c in range(copies) for n, M, K in mats]
    from q4k_b200.shard import shard_rows
    row_range = (lambda M: shard_rows(M, world)[rank]) if sharded else None
    ms = MatSet(torch, q4k_b200, mats, N, device, seed=1234 + (rank if sharded else 0), row_range=row_range)
    total_bytes = sum(algo_bytes(M, K, N) for _, M, K in mats) * (1 if sharded else world)

    gathered = None
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
        gathered = True

        def enqueue():
            for it, sh in zip(ms.items, shards):
                it["full"] = sh.mul_mat(it["x"])
        step = enqueue
        if args.graph and shards[0].gather == "peer":        # NCCL collectives are left out of graph capture
            try:                                               # the whole step (GEMVs + cross-GPU barriers) as one CUDA graph
                step = graph_of(torch, enqueue).replay
                gather_kind += ", CUDA graph"
            except Exception as e:
                torch.cuda.synchronize()
                gather_kind += f", stream launches (graph capture failed: {type(e).__name__})"
    elif args.plan:
        plan = ms.make_plan()                                  # the library's own graph plan (q4k_plan_*)
        step = plan.run
    else:
        g = graph_of(torch, ms.enqueue_device)
        step = g.replay
    n0 = lib.q4k_kernel_launches()
    if plan is not None:
        plan.run()
    else:
        ms.enqueue_device()
    launches_per_step = lib.q4k_kernel_launches() - n0
    torch.cuda.synchronize()

    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ms_total = timed_steps(torch, step, args.steps, args.warmup, barrier)
    # end to end: pinned host activations in, host results out, every step
    if plan is None:
        ms.make_host()
    if sharded and world > 1:
        for it in ms.items:
            it["x_host"] = it["x"].cpu().pin_memory()
            it["full_host"] = torch.zeros((N, it["M_full"]), dtype=torch.float32).pin_memory()
        ms.d2h = sum(it["full_host"].numel() * 4 for it in ms.items)
        def step_host():
            for it, sh in zip(ms.items, shards):
                it["x_stage"].copy_(it["x_host"], non_blocking=True)
                full = sh.mul_mat(it["x_stage"])
                it["full_host"].copy_(full, non_blocking=True)
            torch.cuda.current_stream().synchronize()
    elif plan is not None:
        def step_host():                                       # packed H2D, the graph(s), packed D2H (pipelined inside q4k_plan_run_host)
            plan.run_host(ms.plan_in_host, ms.plan_out_host)
            torch.cuda.current_stream().synchronize()
    else:
        def step_host():
            ms.enqueue_host()
            torch.cuda.current_stream().synchronize()
    ms_e2e = timed_steps(torch, step_host, args.steps, args.warmup, barrier)
    clk = clocks.stop() if rank == 0 else None

    # strong-scaling anchor: the SAME workload, unsharded, on one GPU (every rank measures its own GPU; max over ranks)
    same_1gpu = None
    if sharded and world > 1:
        full = MatSet(torch, q4k_b200, mats, N, device, seed=4321)
        one_plan = full.make_plan()
        ms_one = timed_steps(torch, one_plan.run, args.steps, args.warmup, barrier)
        t1 = torch.tensor([ms_one], device=device, dtype=torch.float64)
        dist.all_reduce(t1, op=dist.ReduceOp.MAX)
        same_1gpu = float(t1[0]) / args.steps

    t = torch.tensor([ms_total, ms_e2e], device=device, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = float(t[0]), float(t[1])
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peak, peak_src, _ = measured_peaks()
    per_step_ms = ms_total / args.steps
    value = total_bytes / (per_step_ms * 1e-3) / 1e9
    e2e_val = total_bytes / (ms_e2e / args.steps * 1e-3) / 1e9
    # dominant kernel = the GEMV: per-launch algorithmic bytes / average launch slot inside the timed region
    n_gemv = int(launches_per_step) if plan is not None else len(ms.items)
    per_launch_bytes = ms.algo_bytes / n_gemv
    per_launch_us = per_step_ms * 1e3 / n_gemv
    achieved = per_launch_bytes / per_launch_us / 1e3
    line = {"metric": "q4k_gemv_hbm_gbs", "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(per_step_ms, 5), "higher_is_better": True,
            "scaling": "strong" if sharded else "weak", "vs_baseline": None,
            "dtype": "int8 x uint4 -> int32, f32 scales", "data": "synthetic (random valid block_q4_K bytes, N(0,1) activations)",
            "config": {"workload": workload_name(args.workload), "batch": N, "mul_mats": len(ms.items),
                       "weight_bytes_per_step_per_gpu": ms.weight_bytes,
                       "l2": f"no flush needed: each step streams {ms.weight_bytes >> 20} MiB of distinct weights per GPU (> 126 MiB L2)"
                             if ms.weight_bytes > L2_BYTES else "WARNING: per-GPU weights fit in L2",
                       "launch": (f"q4k_plan_run: one CUDA graph of {int(launches_per_step)} launches (q/k/v and gate/up fused, shared activations)"
                                  if plan is not None else "CUDA graph replay of the C-ABI launches") if gathered is None else f"gather={gather_kind}",
                       "e2e_path": "q4k_plan_run_host: packed pinned inputs up, four graphs, packed results down; copies pipelined under the kernels on a side stream" if plan is not None
                                   else "host buffers per mul_mat",
                       "parallelism": f"rowshard{world}" if sharded else f"replica{world}"},
            "decode_tok_s": round(1e3 / per_step_ms, 1) if not sharded else None,
            "pct_of_nominal_8TBs": round(100 * value / world / HBM_NOMINAL_GBS, 2),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": NCU_TRAFFIC["bytes"], "traffic_note": NCU_TRAFFIC["note"],
                         "peak_source": peak_src,
                         "kernel": "gemv_q4k_mma_kernel (Q8_K quantizer fused into the launch)",
                         "bytes_per_launch": round(per_launch_bytes), "us_per_launch": round(per_launch_us, 3)},
            "e2e": {"value": round(e2e_val, 2), "unit": "GB/s", "h2d_bytes_per_step": ms.h2d, "d2h_bytes_per_step": ms.d2h,
                    "ms_per_step": round(ms_e2e / args.steps, 5),
                    "decode_tok_s": round(1e3 / (ms_e2e / args.steps), 1) if not sharded else None},
            "gpu_launches": int(launches_per_step * args.steps), "clocks": clk}
    if same_1gpu is not None:
        v1 = total_bytes / (same_1gpu * 1e-3) / 1e9
        line["same_workload_1gpu"] = {"value": round(v1, 2), "unit": "GB/s", "ms_per_step": round(same_1gpu, 5),
                                      "speedup": round(value / v1, 3),
                                      "note": "the same matrices unsharded on ONE GPU (q4k_plan_run), measured in this run; the N = 1 bench line "
                                              "is a different workload (TinyLlama token), so compute strong-scaling efficiency from this field"}
    if world == 1 and not args.no_cpu:
        try:
            line["cpu_baseline"] = cpu_baseline_obj(mats, N, total_bytes)
        except Exception as e:       # the baseline is a reported number, never a dependency of the GPU path
            line["cpu_baseline"] = {"value": None, "unit": "GB/s", "cores": 0, "kind": "reference", "sample": f"unavailable: {e}"}
    if world == 1 and args.shapes:
        line["prefill"] = gemm_table(torch, q4k_b200, device, measured_peaks()[2])
        line["shapes"] = gemv_shape_table(torch, q4k_b200, device,
                                          [(5632, 2048, 1), (4096, 14336, 1), (14336, 4096, 1), (4096, 14336, 8),
                                           (28672, 8192, 1), (8192, 28672, 1), (28672, 8192, 8)])
        line["full_q4_k_m_token"] = full_token_extra(torch, q4k_b200, device, args.steps, args.no_cpu)
        line["q6k_shapes"] = q6k_shape_table(torch, q4k_b200, device, [(2048, 5632, 1), (32000, 2048, 1), (14336, 4096, 1), (28672, 8192, 1)])
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=[None, "tinyllama-token", "ffn70b"])
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--no-graph", dest="graph", action="store_false", help="sharded path: plain stream launches")
    ap.add_argument("--gather", default="auto", choices=["auto", "nccl", "peer"], help="row-shard output exchange (N > 1 GPUs)")
    ap.add_argument("--no-plan", dest="plan", action="store_false", help="single GPU: one q4k_mul_mat per launch instead of the graph plan")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--shapes", action="store_true", default=True, help="add the per-shape GEMV table (C1/C4/C5) and the prefill GEMM figures")
    ap.add_argument("--no-shapes", dest="shapes", action="store_false")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.workload is None:
        args.workload = "tinyllama-token" if max(world, args.gpus) == 1 else "ffn70b"
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, world, local_rank)
        if world > 1:                       # symmetric-memory / NCCL teardown order at interpreter exit is not ours to debug
            sys.stdout.flush()
            os._exit(0)


if __name__ == "__main__":
    main()
