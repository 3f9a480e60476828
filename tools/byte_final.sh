#!/bin/bash
# byte-automaton artefacts of HEAD for profiles/: C3 and C5 at 2 GiB (event and record delivery), C5 at 12.5 GB, launch list
set -u
O=gpurun_out; mkdir -p $O
for w in c3 c5; do
  timeout 600 python bench.py --workload $w --bytes $((2<<30)) --steps 5 --warmup 3 --e2e-steps 2 > $O/${w}_events_2g.json 2> $O/${w}_ev.err; echo "$w events rc=$?"; cut -c1-300 $O/${w}_events_2g.json
  AHA_B200_BYTE_RECORDS=1 timeout 600 python bench.py --workload $w --bytes $((2<<30)) --steps 5 --warmup 3 --no-cpu --e2e-steps 2 > $O/${w}_records_2g.json 2> $O/${w}_rec.err; echo "$w records rc=$?"; cut -c1-300 $O/${w}_records_2g.json
done
timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu --e2e-steps 2 > $O/c5_bench_full.json 2> $O/c5_bench_full.err; echo "c5 full rc=$?"; cut -c1-300 $O/c5_bench_full.json
timeout 900 python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu --e2e-steps 2 > $O/c3_bench_full.json 2> $O/c3_bench_full.err; echo "c3 full rc=$?"; cut -c1-300 $O/c3_bench_full.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/c5_launches.csv \
  python bench.py --workload c5 --bytes $((1<<30)) --steps 1 --warmup 3 --no-cpu --e2e-steps 1 > $O/c5_ncu_l.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"scan_fused_kernel" -s 3 -c 1 \
  -o $O/c5_fused_full -f python bench.py --workload c5 --bytes $((1<<30)) --steps 1 --warmup 3 --no-cpu --e2e-steps 1 > $O/c5_ncu_f.log 2>&1; echo "ncu full rc=$?"
