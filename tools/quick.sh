#!/bin/bash
# quick GPU-box check while tuning: parity tests, then the short bench (prints throughput, step ms, walk ms, e2e, launches)
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 2 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('GB/s %.1f  step %.3f ms  walk %.3f ms  e2e %.1f  launches %d  hits %d' % (d['value'], d['ms_per_step'], d['roofline']['dominant_kernel']['ms'], d['e2e']['value'], d['gpu_launches'], d['config']['hits_per_gpu_step']))"
