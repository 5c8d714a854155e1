#!/bin/bash
# N-GPU pass (gpurun --gpus N): sharded correctness (all exchanges, oracle rows, graph replay) + the bench at N ranks per exchange
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/shard_check.py > gpurun_out/shard_check_$N.log 2>&1; echo "shard_check rc=$?"; grep "SHARD_CHECK\|=False\|Error\|error" gpurun_out/shard_check_$N.log | head -20
for g in ${GATHERS:-ll peer nccl}; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 50 --warmup 5 --gather $g > gpurun_out/bench_${N}gpu_$g.json 2> gpurun_out/bench_${N}gpu_$g.err; echo "bench $g rc=$?"
python -c "import json,sys; d=json.loads(open('gpurun_out/bench_${N}gpu_$g.json').read().strip().splitlines()[-1]); print('$g', d['n_gpus'], 'value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'parity', d['parity_checked'], d['run_info']['launch'])" || tail -5 gpurun_out/bench_${N}gpu_$g.err
done
