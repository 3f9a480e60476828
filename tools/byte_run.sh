#!/bin/bash
# byte-automaton path (C3, C5): parity over every delivery variant, then A/B of event vs record delivery
set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/byte_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/byte_tests.log
for w in c3 c5; do
  timeout 600 python bench.py --workload $w --bytes $((2<<30)) --steps 5 --warmup 3 --no-cpu --e2e-steps 2 > $O/${w}_ev.json 2> $O/${w}_ev.err; echo "$w events rc=$?"; cut -c1-1200 $O/${w}_ev.json; tail -2 $O/${w}_ev.err
  AHA_B200_BYTE_RECORDS=1 timeout 600 python bench.py --workload $w --bytes $((2<<30)) --steps 5 --warmup 3 --no-cpu --e2e-steps 2 > $O/${w}_rec.json 2> $O/${w}_rec.err; echo "$w records rc=$?"; cut -c1-400 $O/${w}_rec.json
done
