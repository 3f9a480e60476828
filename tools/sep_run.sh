#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sep or modes_at_config or char_automaton or quirks" 2>&1 | tail -12
for m in sep; do
timeout 600 python bench.py --mode $m --steps 10 --warmup 3 --e2e-steps 2 > gpurun_out/bench_c2_$m.json 2> gpurun_out/bench_c2_$m.err
tail -3 gpurun_out/bench_c2_$m.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_c2_$m.json').read().strip().splitlines()[-1])
r=d['per_rank'][0]
print('$m GB/s %.1f  step %.3f ms  e2e %.1f  launches %d  hits %d' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['config']['hits_per_gpu_step']))
print('scan', r['scan_ms'], 'walk', r['walk_ms'], 'expand', r['expand_ms'])
print('parity', d['parity'])
PY
done
