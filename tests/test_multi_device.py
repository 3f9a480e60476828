"""aha_b200_init(devices, n) + aha_b200_scan: one process, one host thread on the caller's side, the batch dealt to the
selected GPUs in runs of whole documents, hits landing in document order in one pinned buffer (include/aha_b200.h).
Runs in a process of its own (the device selection is per process and precedes the automata).  With one GPU in the box
the same pipeline is driven over that one device (many small runs); with two or more, over two."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import ctypes as C, json, sys
import numpy as np
sys.path.insert(0, %r)
import torch
import aha_b200
from aha_b200 import _lib, synth
from oracle import oracle as O
n_dev = min(2, torch.cuda.device_count())
L = _lib.lib()
devs = (C.c_int32 * n_dev)(*range(n_dev))
_lib.check(L.aha_b200_init(devs, n_dev))
L.aha_b200_debug_set(3, 1)                      # runs of 1 MB: dozens of runs, dealt round-robin
keys = synth.patterns_zh_ngrams(3000) + synth.patterns_en_words(300)
g, o = aha_b200.AC.compile(keys), O.AC.compile(keys)
rng = np.random.default_rng(3)
lens = rng.integers(0, 300000, 150); lens[::7] = 0
off = np.zeros(len(lens) + 1, np.int64); np.cumsum(lens, out=off[1:])
text = synth.gen_host(synth.MIX, 99, 0, int(off[-1]))
out = {"n_dev": n_dev, "bytes": int(off[-1])}
for mode, sep in [(0, None), (aha_b200.UNITS_CHAR, None), (0, [" "]), (aha_b200.LONGEST, None)]:
    res = g.scan_batch((text, off), mode, sep)
    want_n, want_h = o.scan_mode_mt(text, off, 8, mode, sep)
    hits, dho = res.hits(), res.doc_hit_off()
    # checksum of the per-document streams, as the oracle computes it (bench.py / mix_hit)
    import bench
    t = torch.from_numpy(np.frombuffer(hits.tobytes(), np.int32).reshape(-1, 3).copy()).cuda()
    n, h = bench.gpu_stream_hash(t, torch.from_numpy(dho.copy()).cuda(), len(lens))
    assert (res.n_hits, n, h) == (want_n, want_n, want_h), (mode, sep, res.n_hits, want_n)
    assert dho[0] == 0 and dho[-1] == res.n_hits and (np.diff(dho) >= 0).all()
    out["mode_%%d_%%s" %% (mode, bool(sep))] = int(res.n_hits)
print("RESULT " + json.dumps(out))
'''


def test_scan_deals_runs_to_the_selected_devices():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    p = subprocess.run([sys.executable, "-c", CHILD % ROOT], capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")][-1]
    out = json.loads(line[7:])
    assert out["mode_0_False"] > 0 and out["mode_4_False"] > 0
    print(out)
