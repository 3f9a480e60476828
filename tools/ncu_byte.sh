#!/bin/bash
# ncu --set full of the byte-path kernel on the C3 dictionary (1 GiB of text)
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k "regex:scan_fused_kernel" -s 3 -c 1 -o gpurun_out/r02_c3_fused -f \
    python bench.py --workload c3 --bytes 1073741824 --steps 1 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/r02_c3_ncu.log 2>&1
ncu -i gpurun_out/r02_c3_fused.ncu-rep --page raw --csv > gpurun_out/r02_c3_fused_raw.csv 2>/dev/null
python tools/ncu_keys.py gpurun_out/r02_c3_fused_raw.csv
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02_c3_fused_raw.csv')))
h=rows[0]; r=rows[2]
for k in h:
    if k.startswith('lts__t_sectors.') or k.startswith('lts__t_requests.') or 'lts__t_sectors_srcunit_tex.' in k or k.startswith('lts__throughput') or k.startswith('l1tex__m_xbar2l1tex') or k.startswith('lts__t_sectors_op_read.'):
        print(k, r[h.index(k)], rows[1][h.index(k)])
PY
