"""GPU-box probe for config C4: CedarX#match_longest (ignore_case, sep) over EN text as 4 KiB documents."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import aha_b200
from aha_b200 import synth
torch.cuda.set_device(0)
t = aha_b200.Cedar()
for k in synth.patterns_en_phrases(50000):
    t.insert(k)
ac = aha_b200.AC.compile(t)
D = 4096 // synth.BLOCK * synth.BLOCK
n_docs = (1 << 30) // D
text = synth.gen_device(synth.EN, synth.SEED_TEXT + 3, 0, n_docs * D)
off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
mode = aha_b200.TRIE_LONGEST | aha_b200.IGNORE_CASE
for _ in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = ac.scan_device(text, off, mode, sep=[" "])
    torch.cuda.synchronize()
    print("C4 1 GiB: wall %.2f ms  scan_ms %.2f  hits %d  -> %.1f GB/s" % ((time.perf_counter() - t0) * 1e3, r.scan_ms, r.n_hits, n_docs * D / r.scan_ms / 1e6))
    del r
