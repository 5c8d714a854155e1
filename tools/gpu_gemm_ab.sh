#!/bin/bash
# A/B of prefill GEMM builds (make -C ggml-neon-opt_b200/csrc variants VARIANTS=...): timing per variant
mkdir -p gpurun_out
echo "default"; timeout 120 python tools/gemm_bench.py --iters 20 --shapes 4096x14336x512,14336x4096x512,2048x5632x512 2>&1 | tee gpurun_out/gemm_ab.log
for v in $GEMM_VARIANTS; do
  echo "variant $v"; Q4K_B200_LIB=$PWD/ggml-neon-opt_b200/lib/libq4k_b200_$v.so timeout 120 python tools/gemm_bench.py --iters 20 --shapes 4096x14336x512,14336x4096x512,2048x5632x512 2>&1 | tee -a gpurun_out/gemm_ab.log
done
