#!/usr/bin/env python
"""host-only: builds a workload's dictionary (host cedar + AC.compile + flatten) and prints the blob's size and the time
of each step.  No GPU needed.  usage: tools/blob_stats.py c3|c5|c2|c1"""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from aha_b200 import _lib, ac, synth
from oracle import oracle as O
synth.set_host_generator(lambda corpus, seed, first, n: O.gen_text(corpus, seed, first, n))
name = sys.argv[1]
keys = bench.make_patterns(name)
t0 = time.perf_counter()
da = ac.Cedar()
for k in keys:
    da.insert(k)
t1 = time.perf_counter()
nodes, fails, output, kl = da.compiled_arrays()
t2 = time.perf_counter()
blob, n = C.c_void_p(), C.c_int64()
L = _lib.lib()
_lib.check(L.aha_b200_flatten_host(nodes.ctypes.data, len(fails), fails.ctypes.data, output.ctypes.data,
                                   kl.ctypes.data if len(kl) else None, len(kl), C.byref(blob), C.byref(n)))
t3 = time.perf_counter()
print("%s: %d keys, array_size %d, blob %.1f MB; insert %.2f s, AC.compile %.2f s, flatten %.2f s" % (
    name, len(keys), len(fails), n.value / 1e6, t1 - t0, t2 - t1, t3 - t2))
L.aha_b200_blob_free(blob)
