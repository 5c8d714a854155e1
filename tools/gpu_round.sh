#!/bin/bash
# Round pass on one GPU: tests, smoke, bench (both arms), launch list, full ncu capture of the prefill GEMM, micro rates.
mkdir -p gpurun_out
RUN_REF=1 bash tools/gpu_check.sh
timeout 120 python tools/gemm_bench.py --iters 20 > gpurun_out/gemm_bench.log 2>&1; cat gpurun_out/gemm_bench.log
timeout 60 tools/micro/umma_rate > gpurun_out/umma_rate.txt 2>&1; tail -30 gpurun_out/umma_rate.txt
python bench.py --steps 3 --warmup 3 --no-cpu --no-shapes > gpurun_out/plain_bench.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:q4k -c 900 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-shapes > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
timeout 600 bash tools/gpu_ncu_gemm.sh
ls -la gpurun_out
