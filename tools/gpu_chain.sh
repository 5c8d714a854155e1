#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mma.py tests/test_gpu_plan.py -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
: > gpurun_out/chain.log
SPECS="2048x2048,2048+256+256x2048,5632+5632x2048,2048x5632,5632x2048,14336x4096,28672x8192,8192x28672,4096x28672,2048x28672,1024x28672,4096x14336"
timeout 300 python tools/launch_chain.py --specs $SPECS --tag new >> gpurun_out/chain.log 2>&1
cat gpurun_out/chain.log
python bench.py --no-cpu --no-shapes | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('new bench', d['value'], d['ms_per_step'], d['e2e']['value'])"
