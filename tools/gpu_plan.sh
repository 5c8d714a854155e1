#!/bin/bash
# plan/group check: GPU tests, bench with the plan and without it
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu --no-shapes > gpurun_out/bench_plan.json 2> gpurun_out/bench_plan.err; echo "bench rc=$?"; cat gpurun_out/bench_plan.json; tail -3 gpurun_out/bench_plan.err
timeout 600 python bench.py --no-cpu --no-shapes --no-plan > gpurun_out/bench_noplan.json 2> gpurun_out/bench_noplan.err; echo "bench rc=$?"; cat gpurun_out/bench_noplan.json; tail -3 gpurun_out/bench_noplan.err
