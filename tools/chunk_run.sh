#!/bin/bash
# byte kernel: bytes per lane (AHA_B200_BYTE_CHUNK; smaller chunks leave more of the SM's memory to L1) on C3 / C5, 1 GiB
for w in c3 c5; do for ch in 68 84 100 116 132; do
AHA_B200_BYTE_CHUNK=$ch python tools/ab.py run --workload $w --bytes 1073741824 --steps 5 --reps 1 base 2>&1 | sed "s/^/$w chunk $ch: /"
done; done
