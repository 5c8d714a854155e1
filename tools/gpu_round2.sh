#!/bin/bash
# Round pass 2: all GPU tests, smoke, bench (+ new extras), ncu of the Q6_K kernel and of the big-matrix Q4_K kernel
mkdir -p gpurun_out
bash tools/gpu_check.sh
python tools/ncu_target.py --q6k --M 28672 --K 8192 > gpurun_out/plain_q6.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_q6k -s 2 -c 1 -f -o gpurun_out/prof_q6k_c5 \
    python tools/ncu_target.py --q6k --M 28672 --K 8192 > gpurun_out/ncu_q6.log 2>&1
echo "q6k ncu rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_q6k -s 2 -c 1 -f -o gpurun_out/prof_q6k_down \
    python tools/ncu_target.py --q6k --M 2048 --K 5632 > gpurun_out/ncu_q6b.log 2>&1
echo "q6k down ncu rc=$?"
python tools/ncu_target.py --M 28672 --K 8192 > gpurun_out/plain_big.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_q4k_tma -s 2 -c 1 -f -o gpurun_out/prof_q4k_big_c5 \
    python tools/ncu_target.py --M 28672 --K 8192 > gpurun_out/ncu_big.log 2>&1
echo "big ncu rc=$?"
ls -la gpurun_out | tail -20
