#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_kat_on_gpu_box.py -m gpu -x -q -k "longest or spec_vectors or random_ragged or empty or char_units or modes_at_config or kat or KAT or quirks" 2>&1 | tail -8
for m in longest longest-intersectable; do
timeout 600 python bench.py --mode $m --steps 5 --warmup 3 --e2e-steps 2 > gpurun_out/bench_c2_$m.json 2> gpurun_out/bench_c2_$m.err
tail -3 gpurun_out/bench_c2_$m.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_c2_$m.json').read().strip().splitlines()[-1])
r=d['per_rank'][0]
print('$m GB/s %.1f  step %.3f ms  e2e %.1f  launches %d  hits %d' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['config']['hits_per_gpu_step']))
print('parity', d['parity'])
PY
done
