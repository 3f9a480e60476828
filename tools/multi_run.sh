#!/bin/bash
# N-GPU bench lines on one box: tools/multi_run.sh N [workload] [extra bench args]
N=${1:-8}; W=${2:-c2}; shift 2
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $W --steps 20 --warmup 3 "$@" > gpurun_out/bench_${W}_${N}gpu.json 2> gpurun_out/bench_${W}_${N}gpu.err
tail -2 gpurun_out/bench_${W}_${N}gpu.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_${W}_${N}gpu.json').read().strip().splitlines()[-1])
print('N=%d  %.1f GB/s  step %.3f ms  e2e %.1f  slowest rank %d' % (d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['slowest_rank']))
for r in d['per_rank']:
    print(' rank %d total %.2f scan %.3f/%.3f/%.3f gap %.3f/%.3f/%.3f' % (r['rank'], r['total_ms'], r['scan_ms']['min'], r['scan_ms']['median'], r['scan_ms']['max'], r['host_gap_ms']['min'], r['host_gap_ms']['median'], r['host_gap_ms']['max']))
print('gathered', d.get('gathered'), 'parity', d.get('parity'))
PY
