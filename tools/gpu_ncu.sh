#!/bin/bash
# ncu pass: launch list of the bench command + full capture of the GEMV kernel on two shapes.
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:q4k -c 600 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
python tools/ncu_target.py --M 28672 --K 8192 --N 1 > gpurun_out/plain_c5.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemv -s 2 -c 2 -f -o gpurun_out/prof_c5 \
    python tools/ncu_target.py --M 28672 --K 8192 --N 1 > gpurun_out/ncu_c5.log 2>&1
echo "c5 rc=$?"
python tools/ncu_target.py --M 5632 --K 2048 --N 1 > gpurun_out/plain_c1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemv -s 2 -c 2 -f -o gpurun_out/prof_c1 \
    python tools/ncu_target.py --M 5632 --K 2048 --N 1 > gpurun_out/ncu_c1.log 2>&1
echo "c1 rc=$?"
ls -la gpurun_out
