#!/usr/bin/env python
"""per-call latency of the single-document API form (GpuAC#match(String) -> aha_b200_scan on one short host buffer):
median / p99 of the C call alone (ctypes) for a sentence, a 4 KB and a 64 KB document, C2's dictionary."""
import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aha_b200
from aha_b200 import _lib, synth
from aha_b200._lib import Hits
g = aha_b200.AC.compile(synth.patterns_zh_ngrams(10000))
L = _lib.lib()
text = synth.gen_host(synth.ZH, synth.SEED_TEXT + 1, 0, 1 << 20)
for n in (48, 4096, 65536, 1 << 20):
    for mode, name in ((0, "match(Bytes)"), (aha_b200.UNITS_CHAR, "match(String)")):
        buf = np.ascontiguousarray(text[:n])
        off = np.array([0, n], np.int64)
        ts = []
        for i in range(600):
            ph = C.POINTER(Hits)()
            t0 = time.perf_counter()
            rc = L.aha_b200_scan(g._dfa, buf.ctypes.data, off.ctypes.data, 1, mode, None, 0, C.byref(ph))
            t1 = time.perf_counter()
            assert rc == 0
            nh = ph.contents.n_hits
            L.aha_b200_hits_free(ph)
            if i >= 100:
                ts.append((t1 - t0) * 1e6)
        ts = np.array(ts)
        print("%-14s %8d bytes: median %7.1f us  p99 %7.1f us  (%d hits)" % (name, n, np.median(ts), np.percentile(ts, 99), nh))
