#!/bin/bash
# Q6_K prefill GEMM: parity tests, then timings.  Everything under timeout (a new kernel may hang).
mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests/test_gpu_q6k_gemm.py -x -q 2>&1 | tail -25 > gpurun_out/q6k_gemm_tests.log
cat gpurun_out/q6k_gemm_tests.log
timeout -k 5 300 python tools/q6k_gemm_bench.py > gpurun_out/q6k_gemm_bench.txt 2>&1
cat gpurun_out/q6k_gemm_bench.txt
