#!/bin/bash
# e2e of C4 against the run size of the host pipeline (documents per run)
for r in 32768 65536 131072 262144; do
AHA_B200_TRIE_RUN_DOCS=$r python bench.py --workload c4 --bytes 2147483648 --steps 2 --warmup 3 --no-cpu --e2e-steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('run docs $r: e2e %.1f GB/s' % d['e2e']['value'], d['e2e']['call_ms'])"
done
