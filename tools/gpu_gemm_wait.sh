#!/bin/bash
mkdir -p gpurun_out
for m in 0 1 2; do
  echo "wait_mode=$m"
  Q4K_DBG_GEMM_WAIT=$m python tools/gemm_bench.py --shapes 4096x14336x512,2048x5632x512,14336x4096x512 --iters 20
done 2>&1 | tee gpurun_out/gemm_wait.log
