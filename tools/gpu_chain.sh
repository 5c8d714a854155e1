#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mma.py tests/test_gpu_plan.py -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/chain.log
SPECS="2048x2048,5632+5632x2048,2048x5632,5632x2048,4096x14336,14336x4096,28672x8192,8192x28672"
python tools/launch_chain.py --specs $SPECS --order 2 --tag mma_interleave >> gpurun_out/chain.log 2>&1
Q4K_DBG_MMA_CONTIG=1 python tools/launch_chain.py --specs $SPECS --order 2 --tag mma_contig >> gpurun_out/chain.log 2>&1
cat gpurun_out/chain.log
