#!/bin/bash
# every artefact profiles/ holds for a round, in one GPU call: bench lines, reference arm, ncu launch list, ncu --set full
# usage: tools/bank.sh r02   (files land in gpurun_out/<prefix>_*)
P="${1:-r02}"
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/${P}_c2_bench.json 2> gpurun_out/${P}_c2_bench.err; tail -2 gpurun_out/${P}_c2_bench.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${P}_c2_reference.json 2> gpurun_out/${P}_c2_reference.err
# launch list of the same command (cold-cache, serialised: shares only)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${P}_c2_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/${P}_c2_launches.log 2>&1
timeout 900 bash tools/ncu_full.sh "char_walk_kernel|char_count_kernel|char_expand_kernel" ${P}_c2_full > /dev/null
python tools/ncu_keys.py gpurun_out/${P}_c2_full_raw.csv | head -60
for w in c3 c5; do
  timeout 900 python bench.py --workload $w --bytes 2147483648 --steps 5 --warmup 3 --e2e-steps 2 > gpurun_out/${P}_${w}_bench_2g.json 2> gpurun_out/${P}_${w}_bench_2g.err
done
timeout 900 python bench.py --workload c4 --steps 5 --warmup 3 > gpurun_out/${P}_c4_bench.json 2> gpurun_out/${P}_c4_bench.err
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:scan_fused_kernel" -s 3 -c 1 -o gpurun_out/${P}_c3_fused -f \
    python bench.py --workload c3 --bytes 1073741824 --steps 1 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/${P}_c3_ncu.log 2>&1
ncu -i gpurun_out/${P}_c3_fused.ncu-rep --page raw --csv > gpurun_out/${P}_c3_fused_raw.csv 2>/dev/null
for f in c2_bench c2_reference c3_bench_2g c5_bench_2g c4_bench; do echo "== $f"; tail -c 1500 gpurun_out/${P}_$f.json; echo; done
