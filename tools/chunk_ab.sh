#!/bin/bash
# byte kernel: bytes per lane (warm-up share of the lookups) A/B on C3 and C5, 2 GiB
set -u
O=gpurun_out; mkdir -p $O
for c in 132 164 196; do
  for w in c3 c5; do
    AHA_B200_BYTE_CHUNK=$c timeout 600 python bench.py --workload $w --bytes $((2<<30)) --steps 5 --warmup 3 --no-cpu --e2e-steps 1 > $O/${w}_chunk$c.json 2> $O/${w}_chunk$c.err
    python -c "
import json; d=json.load(open('$O/${w}_chunk$c.json')); print('$w chunk $c:', round(d['value'],1), 'GB/s  step', round(d['ms_per_step'],2), 'kernel', round(d['roofline']['kernel_ms'],2), 'hits', d['config']['hits_per_gpu_step'])"
  done
done
AHA_B200_BYTE_CHUNK=196 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused-byte" > $O/chunk196_tests.log 2>&1; echo "tests(196) rc=$?"; tail -2 $O/chunk196_tests.log
for k in 0 1 0 1; do
  AHA_BENCH_KEEP_PREV=$k python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 > $O/c2_keep$k.json 2>/dev/null
  python -c "
import json; d=json.load(open('$O/c2_keep$k.json')); print('c2 keep_prev=$k:', round(d['value'],1), 'GB/s  step', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms'],3), 'walk', round(d['roofline']['dominant_kernel']['ms'],3))"
done
