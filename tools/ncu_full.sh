#!/bin/bash
# one `ncu --set full` capture of the named kernels of one C2 step (after the program has run once without ncu)
# usage: tools/ncu_full.sh "regex of kernel names" out-prefix [extra bench args]
K="$1"; OUT="$2"; shift 2
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 "$@" > gpurun_out/${OUT}_plain.json 2> gpurun_out/${OUT}_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k "regex:$K" -s 9 -c 3 -o gpurun_out/${OUT} -f \
    python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 "$@" > gpurun_out/${OUT}_ncu.log 2>&1
ncu -i gpurun_out/${OUT}.ncu-rep --page raw --csv > gpurun_out/${OUT}_raw.csv 2>/dev/null
ls -la gpurun_out/${OUT}*
