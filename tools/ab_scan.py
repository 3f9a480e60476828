"""A/B timing helper (GPU box): C2 workload at --bytes, normal vs count-only scan, kernel ms as measured by the library."""
import sys, os, argparse
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import aha_b200
from aha_b200 import synth

ap = argparse.ArgumentParser()
ap.add_argument("--bytes", type=int, default=1 << 30)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
torch.cuda.set_device(0)
keys = synth.patterns_zh_ngrams(10000)
ac = aha_b200.AC.compile(keys)
D = synth.doc_bytes(1 << 20)
n_docs = a.bytes // D
total = n_docs * D
text = synth.gen_device(synth.ZH, synth.SEED_TEXT + 1, 0, total)
off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
for name, mode in (("full", 0), ("count_only", aha_b200.ac.COUNT_ONLY), ("chars", aha_b200.ac.UNITS_CHAR)):
    ms = []
    for _ in range(a.reps + 2):
        r = ac.scan_device(text, off, mode)
        ms.append(r.scan_ms)
    ms = ms[2:]
    print("%-10s %8.3f ms  %7.1f GB/s  hits=%d" % (name, np.median(ms), total / np.median(ms) / 1e6, r.n_hits))
