#!/bin/bash
# fused cross-GPU step (q4k_mul_mat_peers_sync) vs the separate signal kernel: correctness, then the bench at N ranks
N=${1:-2}
mkdir -p gpurun_out
for f in 1 0; do
echo "Q4K_SHARD_FUSED_SYNC=$f"
Q4K_SHARD_FUSED_SYNC=$f timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/shard_check.py 2>&1 | grep "SHARD_CHECK\|bit_equal=False\|Error\|error" | head -5
Q4K_SHARD_FUSED_SYNC=$f timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 50 --warmup 5 --gather auto > gpurun_out/bench_${N}gpu_sync$f.json 2> gpurun_out/bench_${N}gpu_sync$f.err; echo "bench rc=$?"
python -c "import json,sys; d=json.loads(open('gpurun_out/bench_${N}gpu_sync$f.json').read().strip().splitlines()[-1]); print('fused=$f', d['n_gpus'], 'value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])"
done
