#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mma.py tests/test_gpu_parity.py tests/test_gpu_plan.py tests/test_gpu_q6k.py -x -q -m gpu 2>&1 | tail -8 | tee gpurun_out/pytest_big.log
timeout 300 python tools/big_sweep.py 2>&1 | tee gpurun_out/big_sweep.log
