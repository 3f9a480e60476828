#!/bin/bash
# byte kernel: bytes per lane (AHA_B200_BYTE_CHUNK) on C3 / C5, 1 GiB
for w in c3 c5; do for ch in 0 164 196; do
AHA_B200_BYTE_CHUNK=$ch python bench.py --workload $w --bytes 1073741824 --steps 5 --warmup 3 --no-cpu --e2e-steps 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w chunk $ch: %.1f GB/s step %.2f ms' % (d['value'], d['ms_per_step']))"
done; done
