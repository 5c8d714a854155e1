#!/bin/bash
# ncu captures of the batched decode kernel (gemv_q4k_wide.cu): N = 32 at 4096x14336 and N = 8 (Q4K_DBG_WIDE_MIN=1) at 28672x8192
mkdir -p gpurun_out
python tools/ncu_target.py --M 4096 --K 14336 --N 32 > gpurun_out/plain_wide32.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemv_q4k_wide -s 2 -c 1 -f -o gpurun_out/prof_wide_n32 \
    python tools/ncu_target.py --M 4096 --K 14336 --N 32 > gpurun_out/ncu_wide32.log 2>&1
echo "n32 rc=$?"
export Q4K_DBG_WIDE_MIN=1
python tools/ncu_target.py --M 28672 --K 8192 --N 8 > gpurun_out/plain_wide8.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemv_q4k_wide -s 2 -c 1 -f -o gpurun_out/prof_wide_n8 \
    python tools/ncu_target.py --M 28672 --K 8192 --N 8 > gpurun_out/ncu_wide8.log 2>&1
echo "n8 rc=$?"
ls -la gpurun_out/*.ncu-rep
