import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import aha_b200
from aha_b200 import synth
torch.cuda.set_device(0)
ac = aha_b200.AC.compile(synth.patterns_zh_ngrams(10000))
D = synth.doc_bytes(1 << 20)
n_docs = (1 << 30) // D
total = n_docs * D
text = synth.gen_device(synth.ZH, synth.SEED_TEXT + 1, 0, total)
off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
for mode in (0, 1, 1, 1, 0, 1):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = ac.scan_device(text, off, mode)
    torch.cuda.synchronize()
    print("mode", mode, "wall %.2f ms  scan_ms %.2f" % ((time.perf_counter() - t0) * 1e3, r.scan_ms))
    del r
