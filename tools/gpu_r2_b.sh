#!/bin/bash
# round 2, run B: full GPU test suite + bench N = 1 with the batched / streamed decode kernels in the dispatch
mkdir -p gpurun_out
(timeout -k 10 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4) | tee gpurun_out/r2_gputests_c.log
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/r2_bench_1gpu_b.json 2> gpurun_out/r2_bench_1gpu_b.err; echo "bench rc=$?"
tail -c 400 gpurun_out/r2_bench_1gpu_b.err
python - <<EOF
import json
for l in open("gpurun_out/r2_bench_1gpu_b.json"):
    if l.startswith("{"):
        d=json.loads(l)
        print({k:d[k] for k in ("value","ms_per_step","parity_checked")}, d["roofline"]["frac"], d["e2e"]["value"])
        print("token", d["decode_token"]["ms_per_step"], d["decode_token"]["decode_tok_s"])
        for s in d["shapes"]: print({k:s[k] for k in ("M","K","N","us","gbs","frac_measured")})
        for s in d["prefill"]["shapes"]: print({k:s.get(k) for k in ("M","K","N","us","tops","gbs","frac_measured")})
        print(d["prefill"].get("tinyllama_pp512"))
EOF
