#!/bin/bash
# byte path (sparse rows): C3 and C5 dictionaries over 2 GiB of MIX text
mkdir -p gpurun_out
for w in c3 c5; do
  timeout 900 python bench.py --workload $w --bytes 2147483648 --steps 5 --warmup 3 --e2e-steps 2 > gpurun_out/bench_${w}_2g.json 2> gpurun_out/bench_${w}_2g.err
  tail -2 gpurun_out/bench_${w}_2g.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_${w}_2g.json').read().strip().splitlines()[-1])
r=d['per_rank'][0]
print('$w GB/s %.1f  step %.3f ms  frac %.3f  e2e %.1f  hits %d  build %.1fs states %d' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['config']['hits_per_gpu_step'], d['dfa_build_s'], d['config']['dfa_states']))
print('scan', r['scan_ms'], 'parity', d['parity'], 'cpu', d['cpu_baseline']['value'])
PY
done
