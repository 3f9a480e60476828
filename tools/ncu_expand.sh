#!/bin/bash
# --set full + source page of char_count_kernel / char_expand_kernel of one C2 step
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k "regex:char_count_kernel|char_expand_kernel" -s 6 -c 2 -o gpurun_out/r02_c2_expand -f \
    python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/r02_c2_expand_ncu.log 2>&1
ncu -i gpurun_out/r02_c2_expand.ncu-rep --page raw --csv > gpurun_out/r02_c2_expand_raw.csv 2>/dev/null
ncu -i gpurun_out/r02_c2_expand.ncu-rep --page source --csv --kernel-name regex:char_expand_kernel > gpurun_out/r02_c2_expand_source.csv 2>/dev/null
ncu -i gpurun_out/r02_c2_expand.ncu-rep --page source --csv --kernel-name regex:char_count_kernel > gpurun_out/r02_c2_count_source.csv 2>/dev/null
python tools/ncu_keys.py gpurun_out/r02_c2_expand_raw.csv | grep -E "=====|time_duration|inst_executed.sum|issue_active|lsu_wavefronts.avg|thread_inst"
