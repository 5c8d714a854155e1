#!/bin/bash
# round 2, run A: plugin tests (cached graph / plans / host buffer type), bench N = 1, reference arm, test-backend-ops perf
mkdir -p gpurun_out
python -m pytest tests/test_plugin.py tests/test_gpu_plan.py -m gpu -q 2>&1 | tail -15 > gpurun_out/r2_plugin_tests.log
tail -3 gpurun_out/r2_plugin_tests.log
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/r2_bench_1gpu_a.json 2> gpurun_out/r2_bench_1gpu_a.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2_bench_1gpu_a.err
timeout 300 python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/r2_bench_reference_a.json 2>&1; echo "ref rc=$?"
export GGML_BACKEND_PATH=$PWD/ggml-neon-opt_b200/lib/libggml-q4k-b200.so
timeout 600 oracle/_ref/test-backend-ops perf -o MUL_MAT -b Q4K_B200_0 -p type_a=q4_K > gpurun_out/r2_test_backend_ops_perf_q4k.txt 2>&1; echo "tbo perf rc=$?"
tail -25 gpurun_out/r2_test_backend_ops_perf_q4k.txt
