#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "gemm" 2>&1 | tail -4 | tee gpurun_out/pytest_split.log
for s in 0 1 2 4; do
echo "--- KSPLIT=$s (0 = heuristic)"
if [ $s = 0 ]; then unset Q4K_DBG_GEMM_KSPLIT; else export Q4K_DBG_GEMM_KSPLIT=$s; fi
timeout 200 python tools/gemm_bench.py --iters 20 --shapes 5632x2048x512,2048x5632x512,2048x2048x512,256x2048x512,4096x14336x32,14336x4096x32,4096x4096x128 2>&1 | tee -a gpurun_out/gemm_bench_split.log
done
