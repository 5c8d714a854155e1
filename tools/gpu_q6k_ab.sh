#!/bin/bash
# A/B of Q6_K GEMM builds: libq4k_b200_<variant>.so next to the default build
mkdir -p gpurun_out
: > gpurun_out/q6k_ab.txt
for v in default $Q6K_VARIANTS; do
  echo "== $v" | tee -a gpurun_out/q6k_ab.txt
  if [ $v = default ]; then L=; else L=$PWD/ggml-neon-opt_b200/lib/libq4k_b200_$v.so; fi
  Q4K_B200_LIB=$L timeout -k 5 200 python tools/q6k_gemm_bench.py --gemm-only $Q6K_BENCH_ARGS 2>&1 | tee -a gpurun_out/q6k_ab.txt
done
