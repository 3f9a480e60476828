#!/bin/bash
# byte kernel in span mode (AHA_B200_BYTE_SPAN bytes per lane, no staging) against the staged tiles: C3 / C5, 1 GiB
# usage: tools/span_run.sh "<so tags>" "<spans>"
for w in c3 c5; do for tag in $1; do for sp in $2; do
AHA_B200_SO=aha_b200/libaha_b200_$tag.so AHA_B200_BYTE_SPAN=$sp python bench.py --workload $w --bytes 1073741824 --steps 5 --warmup 3 --e2e-steps 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w $tag span $sp: %.1f GB/s step %.2f ms parity %s' % (d['value'], d['ms_per_step'], d['parity']['equal']))"
done; done; done
