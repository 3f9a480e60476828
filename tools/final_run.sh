#!/bin/bash
# One GPU-box pass that produces every artefact profiles/ cites for HEAD (run under gpurun, one GPU):
#   tests, smoke, the bench line, the reference arm, the ncu launch list, one `--set full` capture of the
#   two dominant kernels at the bench workload.  Numbers printed by the runs under ncu are not bench values.
set -u
TAG=${1:-r01_final}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/${TAG}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${TAG}_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; cat $O/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_reference.json 2>> $O/${TAG}_bench.err; echo "ref rc=$?"; cat $O/${TAG}_reference.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 1 > $O/${TAG}_ncu_l.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"char_walk_kernel|char_expand_kernel" -s 6 -c 2 \
  -o $O/${TAG}_walk_expand_full -f python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 1 > $O/${TAG}_ncu_f.log 2>&1; echo "ncu full rc=$?"
ls -la $O | grep ${TAG}
