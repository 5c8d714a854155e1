#!/bin/bash
# Full single-GPU pass: tests, smoke, bench (both arms), ncu launch list of the bench command.
mkdir -p gpurun_out
bash tools/gpu_check.sh
python bench.py --steps 3 --warmup 3 --no-cpu --no-shapes > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:q4k -c 900 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-shapes > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
