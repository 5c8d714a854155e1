#!/bin/bash
# development loop of the prefill GEMM: parity tests, timing
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "gemm" 2>&1 | tail -4
timeout 120 python tools/gemm_bench.py --iters 20 2>&1 | tee gpurun_out/gemm_bench.log
