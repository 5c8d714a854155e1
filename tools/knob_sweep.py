"""One GPU: microseconds per q4k_mul_mat launch for a list of shapes under a list of knob settings (q4k_debug_set), launches
captured in a CUDA graph and rotated through > 2.5 x L2 of distinct weight copies.  Development tool.
    python tools/knob_sweep.py --shapes 4096x28672x1,1024x28672x1 --sets "auto:;kp2:mma_kparts=2;big:big_matrix_bytes=1"
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-neon-opt_b200"))
import q4k_b200  # noqa: E402

HBM = 6538.3


def rand_q4k(M, K, copies):
    nb = K // 256
    w = torch.randint(0, 256, (copies, M, nb, 144), dtype=torch.uint8, device="cuda")
    w[..., 0] = 0x00; w[..., 1] = 0x2c; w[..., 2] = 0x00; w[..., 3] = 0x28
    return w.view(copies, M, nb * 144)


def graph_time_us(launch, per_graph, replays=10):
    launch(0)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(per_graph):
            launch(i)
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(True), torch.cuda.Event(True)
    s.record()
    for _ in range(replays):
        g.replay()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) * 1e3 / (per_graph * replays)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", required=True)
    ap.add_argument("--sets", default="auto:")
    a = ap.parse_args()
    sets = []
    for s in a.sets.split(";"):
        name, _, kv = s.partition(":")
        sets.append((name, [tuple(x.split("=")) for x in kv.split(",") if x]))
    for shp in a.shapes.split(","):
        M, K, N = map(int, shp.split("x"))
        wbytes = M * (K // 256) * 144
        copies = max(2, (320 << 20) // wbytes + 1)
        w = rand_q4k(M, K, copies)
        x = torch.randn((N, K), device="cuda")
        dst = torch.empty((N, M), device="cuda")
        ws = torch.empty(max(q4k_b200.workspace_bytes(K, N), 16), dtype=torch.uint8, device="cuda")
        row = {"M": M, "K": K, "N": N, "MB": round(wbytes / 1e6, 1), "hbm_floor_us": round((wbytes + N * K * 4 + N * M * 4) / HBM / 1e3, 2)}
        for name, kvs in sets:
            q4k_b200.debug_reset()
            for k, v in kvs:
                q4k_b200.debug_set(k, int(v))
            try:
                n0 = q4k_b200.kernel_launches()
                q4k_b200.mul_mat(w[0], x, dst, ws)
                nl = q4k_b200.kernel_launches() - n0
                us = graph_time_us(lambda i: q4k_b200.mul_mat(w[i % copies], x, dst, ws), per_graph=max(copies, 20))
                row[name] = round(us, 2)
                row[name + "_launches"] = nl
            except Exception as e:
                row[name] = f"{type(e).__name__}"
        q4k_b200.debug_reset()
        print(json.dumps(row), flush=True)
        del w
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
