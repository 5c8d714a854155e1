#!/bin/bash
# Q6_K development pass: parity tests, sanitizer on the unaligned-tail test, sweep at two ring depths
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_q6k.py -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/pytest_q6k.log
timeout 300 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_q6k.py -x -q -m gpu -k "unaligned or zero_act or golden" 2>&1 | tail -8 | tee gpurun_out/sanitizer_q6k.log
timeout 300 python tools/q6k_sweep.py --out gpurun_out/q6k_sweep.json 2>&1 | tee gpurun_out/q6k_sweep.log
for s in 1 3 4; do echo "stages $s"; Q6K_DBG_STAGES=$s timeout 200 python tools/q6k_sweep.py --ns 1 --shapes 2048x5632,32000x2048,28672x8192 2>&1 | tee -a gpurun_out/q6k_sweep_stages.log; done
