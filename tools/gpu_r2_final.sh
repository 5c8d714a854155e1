#!/bin/bash
# round 2 final pass on one GPU: tests, smoke, bench (both arms), ncu launch list + full captures of the headline kernels,
# the reference's test-backend-ops perf through the plugin, llama-bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
(timeout -k 10 900 python -m pytest tests -m gpu -q 2>&1 | tail -4) | tee gpurun_out/r2_gputests_final.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke.log
timeout 900 python bench.py > gpurun_out/r2_bench_1gpu_final.json 2> gpurun_out/r2_bench_1gpu_final.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/r2_bench_reference_final.json 2>/dev/null; echo "ref rc=$?"
python bench.py --steps 3 --warmup 3 --no-cpu --no-shapes > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-shapes > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
python tools/ncu_target.py --M 28672 --K 8192 --N 1 > gpurun_out/plain_c5a.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemv_q4k -s 2 -c 1 -f -o gpurun_out/prof_r2_c5_28672x8192 \
    python tools/ncu_target.py --M 28672 --K 8192 --N 1 > gpurun_out/ncu_c5a.log 2>&1
echo "c5a rc=$?"
python tools/ncu_target.py --M 8192 --K 28672 --N 1 > gpurun_out/plain_c5b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemv_q4k -s 2 -c 1 -f -o gpurun_out/prof_r2_c5_8192x28672 \
    python tools/ncu_target.py --M 8192 --K 28672 --N 1 > gpurun_out/ncu_c5b.log 2>&1
echo "c5b rc=$?"
timeout -k 5 300 python tools/q6k_gemm_bench.py > gpurun_out/r2_q6k_gemm_bench.txt 2>&1; echo "q6k bench rc=$?"; head -4 gpurun_out/r2_q6k_gemm_bench.txt
timeout -k 5 120 python tools/q6k_gemm_one.py 4096 14336 512 > gpurun_out/plain_q6k.log 2>&1 &&
timeout -k 5 400 ncu --set full --clock-control none --import-source on -k regex:gemm_q6k_tc -s 2 -c 1 -f -o gpurun_out/prof_r2_q6k_gemm_4096x14336x512 \
    python tools/q6k_gemm_one.py 4096 14336 512 > gpurun_out/ncu_q6k.log 2>&1
echo "q6k ncu rc=$?"
export GGML_BACKEND_PATH=$PWD/ggml-neon-opt_b200/lib/libggml-q4k-b200.so
timeout 600 oracle/_ref/test-backend-ops perf -o MUL_MAT -b Q4K_B200_0 -p type_a=q4_K > gpurun_out/r2_test_backend_ops_perf_q4k_final.txt 2>&1; echo "tbo perf rc=$?"
unset GGML_BACKEND_PATH
bash tools/gpu_llama_bench.sh > /dev/null 2>&1; tail -12 gpurun_out/r2_llama_bench.txt
python - <<EOF
import json
for l in open("gpurun_out/r2_bench_1gpu_final.json"):
    if l.startswith("{"):
        d=json.loads(l)
        print({k:d[k] for k in ("value","ms_per_step","parity_checked")}, d["roofline"]["frac"], d["e2e"]["value"], d["clocks"])
        print("token", d["decode_token"]["ms_per_step"], d["decode_token"]["e2e"], d["decode_token"].get("e2e_plugin"))
        print(d["prefill"].get("tinyllama_pp512")); print(d.get("llama_bench")); print(d["cpu_baseline"]["value"])
EOF
