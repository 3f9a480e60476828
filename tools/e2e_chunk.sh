#!/bin/bash
# e2e of C2 against the run size of the host pipeline
for mb in 64 64 32; do
AHA_B200_E2E_CHUNK_MB=$mb python bench.py --steps 3 --warmup 3 --no-cpu --e2e-steps 8 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('run MB $mb: e2e %.1f GB/s' % d['e2e']['value'], d['e2e']['call_ms'])"
done
