#!/bin/bash
# Round-end pass on one GPU: all tests, smoke, bench (both arms), ncu launch list of the bench command
mkdir -p gpurun_out
RUN_REF=1 bash tools/gpu_check.sh
python bench.py --steps 3 --warmup 3 --no-cpu --no-shapes > gpurun_out/plain_bench.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:q4k\|q6k -c 900 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-shapes > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
timeout 200 python tools/q6k_sweep.py --ns 1 2>&1 | tee gpurun_out/q6k_sweep3.log
