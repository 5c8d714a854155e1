#!/bin/bash
mkdir -p gpurun_out
python tools/gemm_bench.py --shapes 4096x14336x512 --iters 3 > gpurun_out/plain_gemm.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_q4k_tc -s 2 -c 1 -f -o gpurun_out/prof_gemm \
    python tools/gemm_bench.py --shapes 4096x14336x512 --iters 3 > gpurun_out/ncu_gemm.log 2>&1
echo "gemm rc=$?"
