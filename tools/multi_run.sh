#!/bin/bash
# 8-GPU lines for the configurations BASELINE quotes on 8 GPUs: C5 (1 M patterns, 100 GB = 8 x 12.5 GB, blob by ncclBroadcast), C3.
set -u
N=${1:-8}
O=gpurun_out; mkdir -p $O
for w in c5 c3; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --workload $w --steps 3 --warmup 3 --no-cpu --e2e-steps 2 > $O/${w}_bench_${N}gpu.json 2> $O/${w}_bench_${N}gpu.err
  echo "$w x$N rc=$?"; cat $O/${w}_bench_${N}gpu.json; tail -2 $O/${w}_bench_${N}gpu.err
done
