#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_target.py --M 28672 --K 8192 --N 8 > gpurun_out/plain_n8.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_q4k_mma -s 2 -c 1 -f -o gpurun_out/prof_mma_n8_c5 \
    python tools/ncu_target.py --M 28672 --K 8192 --N 8 > gpurun_out/ncu_n8.log 2>&1
echo "n8 rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_q4k_mma -s 2 -c 1 -f -o gpurun_out/prof_mma_n4_c5 \
    python tools/ncu_target.py --M 28672 --K 8192 --N 4 > gpurun_out/ncu_n4.log 2>&1
echo "n4 rc=$?"
