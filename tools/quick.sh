#!/bin/bash
# quick GPU-box check while tuning: parity tests, then the bench line (saved under gpurun_out/)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
tail -3 gpurun_out/bench_c2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_c2.json').read().strip().splitlines()[-1])
r=d['per_rank'][0]
print('GB/s %.1f  step %.3f ms  frac %.3f  e2e %.1f  launches %d  hits %d' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['gpu_launches'], d['config']['hits_per_gpu_step']))
print('scan', r['scan_ms'], 'walk', r['walk_ms'], 'expand', r['expand_ms'], 'gap', r['host_gap_ms'])
print('parity', d['parity'])
print('cpu', d['cpu_baseline'])
PY
