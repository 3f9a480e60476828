#!/bin/bash
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k "regex:trie_folded_doc_kernel" -s 6 -c 1 -o gpurun_out/r02_c4_trie -f \
    python bench.py --workload c4 --bytes 1073741824 --steps 1 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/r02_c4_ncu.log 2>&1
ncu -i gpurun_out/r02_c4_trie.ncu-rep --page raw --csv > gpurun_out/r02_c4_trie_raw.csv 2>/dev/null
python tools/ncu_keys.py gpurun_out/r02_c4_trie_raw.csv
ncu -i gpurun_out/r02_c4_trie.ncu-rep --page source --csv > gpurun_out/r02_c4_trie_source.csv 2>/dev/null
ls -la gpurun_out/r02_c4*
