#!/bin/bash
# N-GPU pass (gpurun --gpus N): sharded correctness + the bench at N ranks
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/shard_check.py > gpurun_out/shard_check_$N.log 2>&1; echo "shard_check rc=$?"; grep -v "^W\|^\[W\|Warning" gpurun_out/shard_check_$N.log | tail -12
for g in nccl auto; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 50 --warmup 5 --gather $g > gpurun_out/bench_${N}gpu_$g.json 2> gpurun_out/bench_${N}gpu_$g.err; echo "bench $g rc=$?"; cat gpurun_out/bench_${N}gpu_$g.json; tail -3 gpurun_out/bench_${N}gpu_$g.err
done
