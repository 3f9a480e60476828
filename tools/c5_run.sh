#!/bin/bash
# C5 (1 M patterns): scaled parity test, bench lines at 2 GiB and at the full 12.5 GB per GPU, launch list.
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k c5 > $O/c5_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/c5_tests.log
timeout 900 python bench.py --workload c5 --bytes $((2<<30)) --steps 3 --warmup 3 > $O/c5_bench_2g.json 2> $O/c5_bench_2g.err; echo "bench 2g rc=$?"; cat $O/c5_bench_2g.json; tail -3 $O/c5_bench_2g.err
timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu --e2e-steps 2 > $O/c5_bench_full.json 2> $O/c5_bench_full.err; echo "bench full rc=$?"; cat $O/c5_bench_full.json; tail -3 $O/c5_bench_full.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/c5_launches.csv \
  python bench.py --workload c5 --bytes $((1<<30)) --steps 1 --warmup 3 --no-cpu --e2e-steps 1 > $O/c5_ncu_l.log 2>&1; echo "ncu list rc=$?"
