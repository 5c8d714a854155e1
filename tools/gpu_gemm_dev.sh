#!/bin/bash
# development loop of the prefill GEMM: parity tests, timing, per-role trace
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "gemm" 2>&1 | tail -15
timeout 120 python tools/gemm_bench.py --iters 20 2>&1 | tee gpurun_out/gemm_bench.log
for v in $GEMM_VARIANTS; do
  echo "variant $v"; Q4K_B200_LIB=$PWD/ggml-neon-opt_b200/lib/libq4k_b200_$v.so timeout 120 python tools/gemm_bench.py --iters 20 --shapes 4096x14336x512,2048x5632x512 2>&1 | tee -a gpurun_out/gemm_bench.log
done
timeout 120 python tools/gemm_trace.py 2>&1 | grep -v "^s " | tee gpurun_out/gemm_trace.log
