#!/usr/bin/env python
"""host-only: which states of a workload's byte automaton the walk visits (sparse rows, no GPU): share of the lookups
that fall on wide states and on the first K narrow states (BFS order), and iterations per byte.
usage: tools/visit_stats.py c3|c5 [sample_bytes]"""
import ctypes as C, os, struct, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from aha_b200 import _lib, ac, synth
from oracle import oracle as O
synth.set_host_generator(lambda corpus, seed, first, n: O.gen_text(corpus, seed, first, n))
name = sys.argv[1]
nbytes = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
w = bench.WORKLOADS[name]
keys = bench.make_patterns(name)
da = ac.Cedar()
for k in keys:
    da.insert(k)
nodes, fails, output, kl = da.compiled_arrays()
blob, n = C.c_void_p(), C.c_int64()
L = _lib.lib()
_lib.check(L.aha_b200_flatten_host(nodes.ctypes.data, len(fails), fails.ctypes.data, output.ctypes.data,
                                   kl.ctypes.data if len(kl) else None, len(kl), C.byref(blob), C.byref(n)))
raw = C.string_at(blob.value, n.value)
L.aha_b200_blob_free(blob)
(magic, ver, Cc, S, K, n_out, maxlen, _r, o_cmap, o_delta, o_oo, o_ov, o_kl, o_kt, o_dep, o_fail, total, o_rows, o_wide, n_wide,
 has_delta, wide_first, o_wf) = struct.unpack_from("<QIIQQQII" + "Q" * 15, raw, 0)
print("states %d classes %d wide %d (%.1f KB of wide rows) has_delta %d" % (S, Cc, n_wide, n_wide * Cc * 4 / 1024, has_delta))
cmap = np.frombuffer(raw, np.uint8, 256, o_cmap)
rows = np.frombuffer(raw, "<u4", S * 4, o_rows).reshape(S, 4)
wide = np.frombuffer(raw, "<u4", n_wide * Cc, o_wide).reshape(n_wide, Cc)
text = O.gen_text(bench.CORPUS_ID[w["corpus"]], bench.SEED_TEXT + w["seed"], 0, nbytes)
visits = np.zeros(S, np.int64)
rows_l, wide_l, cmap_l = rows.tolist(), wide.tolist(), cmap.tolist()
s, iters = 0, 0
for byte in text.tolist():
    while True:
        visits[s] += 1
        iters += 1
        if s < n_wide and wide_first:
            e = wide_l[s][cmap_l[byte]]
            s = e >> 9
            break
        r = rows_l[s]
        e = 0
        for k in (1, 2, 3):
            if r[k] and (r[k] & 0xFF) == byte:
                e = r[k]
        if e:
            s = e >> 9
            break
        s = r[0] >> 8
print("iterations per byte %.3f" % (iters / nbytes))
tot = visits.sum()
print("lookups on wide states: %.3f" % (visits[:n_wide].sum() / tot))
dep = np.frombuffer(raw, "<u2", S, o_dep)
for d in range(0, 8):
    print("  depth %d: %.3f of the lookups, %d states" % (d, visits[dep == d].sum() / tot, int((dep == d).sum())))
order = np.argsort(-visits)
cum = np.cumsum(visits[order]) / tot
for k in (64, 160, 512, 1024, 2048, 4096, 8192, 16384, 65536):
    if k <= S:
        sel = order[:k]
        nw = int((sel < n_wide).sum())
        print("  hottest %6d states: %.3f of the lookups (%d wide, %d narrow: %.0f KB)" % (k, cum[k - 1], nw, k - nw, (nw * Cc * 4 + (k - nw) * 16) / 1024))
