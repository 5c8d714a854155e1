#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
for b in 0 1; do
echo "BIG_BYTES=$b"
BIG_BYTES=$b timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/shard_timing.py 2>&1 | grep "M=" | tee -a gpurun_out/shard_ab_$N.log
done
