#!/bin/bash
# per-phase warp cycles of the byte kernel (build with -DAHA_B200_PHASES: tools/ab.py build ph:-DAHA_B200_PHASES)
for w in c3 c5; do
AHA_B200_SO=aha_b200/libaha_b200_ph.so python bench.py --workload $w --bytes 1073741824 --steps 3 --warmup 3 --no-cpu --e2e-steps 0 2>&1 >/dev/null | grep phases | tail -2
done
