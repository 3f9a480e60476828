#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "cedar or c4 or trie" 2>&1 | tail -4
timeout 900 python bench.py --workload c4 --steps 5 --warmup 3 ${C4_ARGS} > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
tail -3 gpurun_out/bench_c4.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_c4.json').read().strip().splitlines()[-1])
r=d['per_rank'][0]
print('GB/s %.1f  step %.3f ms  frac %.3f  e2e %.1f  launches %d  hits %d' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['gpu_launches'], d['config']['hits_per_gpu_step']))
print('scan', r['scan_ms'], 'gap', r['host_gap_ms'])
print('parity', d['parity'])
print('cpu', d['cpu_baseline'])
PY
