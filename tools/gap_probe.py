"""Where does the time between the library's scan events (scan_ms) and the wall time of a scan_device call go?
usage: gap_probe.py [workload] [bytes]   -- prints, per call: wall ms of the call, scan_ms, wall ms of releasing the previous result."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import aha_b200
from aha_b200 import synth
import bench
name = sys.argv[1] if len(sys.argv) > 1 else "c5"
nbytes = int(sys.argv[2]) if len(sys.argv) > 2 else 2 << 30
keys, corpus, seed = bench.make_patterns(name)
ac = aha_b200.AC.compile(keys)
D = synth.doc_bytes(1 << 20)
n_docs = nbytes // D
text = synth.gen_device(corpus, seed, 0, n_docs * D)
doc_off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
res = None
for i in range(8):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = None  # release the previous result first (as a caller with one result in flight does)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    res = ac.scan_device(text, doc_off)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("call %d: release %.2f ms  scan_device %.2f ms  scan_ms %.2f  hits %d  mem reserved %.1f GB" % (
        i, (t1 - t0) * 1e3, (t2 - t1) * 1e3, res.scan_ms, res.n_hits, (torch.cuda.mem_get_info()[1] - torch.cuda.mem_get_info()[0]) / 1e9), flush=True)
# and with the previous result still alive while the next is produced (what bench.py's loop does)
for i in range(4):
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    res = ac.scan_device(text, doc_off)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("overlapped %d: scan_device %.2f ms  scan_ms %.2f" % (i, (t2 - t1) * 1e3, res.scan_ms), flush=True)
