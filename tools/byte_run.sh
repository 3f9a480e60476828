#!/bin/bash
# byte-automaton path (C3, C5): parity over every delivery variant, then A/B of event vs record delivery
set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/byte_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/byte_tests.log
for w in c3 c5; do
  timeout 600 python bench.py --workload $w --bytes $((2<<30)) --steps 5 --warmup 3 --no-cpu --e2e-steps 2 > $O/${w}_ev.json 2> $O/${w}_ev.err; echo "$w events rc=$?"; cut -c1-1200 $O/${w}_ev.json; tail -2 $O/${w}_ev.err
  AHA_B200_BYTE_RECORDS=1 timeout 600 python bench.py --workload $w --bytes $((2<<30)) --steps 5 --warmup 3 --no-cpu --e2e-steps 2 > $O/${w}_rec.json 2> $O/${w}_rec.err; echo "$w records rc=$?"; cut -c1-400 $O/${w}_rec.json
done
# host-side stamps of one call (where the time between the kernel and the step goes): phases build, last because it rebuilds the .so
AHA_B200_NVCC_EXTRA=-DAHA_B200_PHASES AHA_FORCE_BUILD=1 python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
python bench.py --workload c5 --bytes $((2<<30)) --steps 2 --warmup 3 --no-cpu --e2e-steps 1 2> $O/c5_phases.err | cut -c1-200; grep -E "^\[host\]|^\[phases\]" $O/c5_phases.err | tail -24
