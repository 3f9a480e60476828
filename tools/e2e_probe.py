"""GPU box: time aha_b200_scan (host buffers) for several pipeline run sizes."""
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
dev = synth.gen_device(synth.ZH, synth.SEED_TEXT + 1, 0, total)
host = torch.empty(total, dtype=torch.uint8, pin_memory=True)
host.copy_(dev)
del dev
hn = host.numpy()
off = np.arange(n_docs + 1, dtype=np.int64) * D
for mb in (int(x) for x in (sys.argv[1:] or ["16", "64", "256", "1100"])):
    os.environ["AHA_B200_E2E_CHUNK_MB"] = str(mb)
    ts = []
    for _ in range(4):
        t0 = time.perf_counter()
        r = ac.scan_batch((hn, off))
        n = r.n_hits
        ts.append(time.perf_counter() - t0)
        sm = r.scan_ms
        del r
    print("run=%4d MB  best %.1f ms  %.1f GB/s  (kernels %.1f ms, hits %d)" % (mb, min(ts[1:]) * 1e3, total / min(ts[1:]) / 1e9, sm, n))
