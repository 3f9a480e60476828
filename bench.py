#!/usr/bin/env python
"""bench.py -- AC scan GB/s of text (+ hits/s) on N x B200, beside Aha::AC on the host cores.

Default workload (BASELINE.json configs[1], "C2"): 10k CJK UTF-8 patterns over 2 GB of synthetic Chinese text per
GPU (2048 documents of ~1 MiB), dictionary sampled from the same stream.  One step = one pass of the scan over the
whole batch.  `value` times the scan with text resident in HBM; `e2e` times the C-ABI call aha_b200_scan() on HOST
buffers (H2D text, scan, D2H hits inside the timed region).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c1|c2|c3|c4|c5] [--gather]
N > 1 is launched by torchrun (one rank per GPU): the corpus shards by document (weak scaling), the DFA is built on
rank 0 and broadcast over NCCL, hit counts are all-gathered.  Per-step timings of every rank (library events and the
host gap around them) are gathered and printed in `per_rank`, so a slow step or a slow rank is visible.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

UNITS_CHAR, SEP, LONGEST, INTERSECTABLE, TRIE_LONGEST, IGNORE_CASE = 1, 2, 4, 8, 16, 32

WORKLOADS = {
    # name: corpus, seed offset, patterns, bytes per GPU, doc bytes, mode, sep, description
    "c1": dict(corpus="EN", seed=0, n_pat=1000, nbytes=100 << 20, doc=1 << 20, mode=0, sep=None,
               desc="C1: 1k ASCII word patterns / 100 MB EN text"),
    "c2": dict(corpus="ZH", seed=1, n_pat=10000, nbytes=2 << 30, doc=1 << 20, mode=0, sep=None,
               desc="C2: 10k CJK patterns / 2 GB ZH text"),
    "c3": dict(corpus="MIX", seed=2, n_pat=100000, nbytes=10 << 30, doc=1 << 20, mode=0, sep=None,
               desc="C3: 100k patterns / 10 GB MIX text"),
    # configs[3]: CedarX#match_longest(ignore_case, sep = [' ']) over 5 GB of EN text as 4 KiB documents
    "c4": dict(corpus="EN", seed=3, n_pat=50000, nbytes=5 << 30, doc=4096, mode=TRIE_LONGEST | IGNORE_CASE, sep=[" "],
               desc="C4: CedarX#match_longest, 50k mixed-case EN words and phrases, ignore_case, sep=[' '] / 5 GB EN text as 4 KiB documents"),
    # configs[4]: 100 GB over 8 GPUs = 12.5 GB per GPU, the byte automaton broadcast over NCCL
    "c5": dict(corpus="MIX", seed=2, n_pat=1000000, nbytes=12500 << 20, doc=1 << 20, mode=0, sep=None,
               desc="C5: 1M patterns / 100 GB MIX text document-sharded over 8 GPUs (12.5 GB per GPU)"),
}
CORPUS_ID = {"EN": 0, "ZH": 1, "MIX": 2}
SEED_TEXT = 0xA11A0002

# order-sensitive stream checksum shared with oracle/aha_oracle.cpp (mix_hit): constants as signed 64-bit
_M64 = (1 << 64) - 1


def _s64(x):
    x &= _M64
    return x - (1 << 64) if x >> 63 else x


_KA, _KB, _KC, _KD, _KE, _KF = (_s64(v) for v in (0x9E3779B97F4A7C15, 0xC2B2AE3D27D4EB4F, 0x165667B19E3779F9,
                                                  0x85EBCA77C2B2AE63, 0x27D4EB2F165667C5, 0xBF58476D1CE4E5B9))


def make_patterns(name):
    from aha_b200 import synth
    w = WORKLOADS[name]
    if name == "c1":
        return synth.patterns_en_words(1000)
    if name == "c2":
        return synth.patterns_zh_ngrams(10000)
    if name == "c4":
        return synth.patterns_en_phrases(w["n_pat"])
    return synth.patterns_mix(w["n_pat"])


def measured_traffic(name):
    """dram__bytes_read.sum + dram__bytes_write.sum of the scan kernels from the committed `ncu --set full` capture
    (profiles/traffic.json: bytes per text byte for the workload), scaled to this launch by the caller."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(name)
    except Exception:
        return None


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---- CPU side: the oracle port of the reference (test infrastructure; the only place bench.py executes oracle/) ----
class CpuSide:
    """The reference's CPU path for the workload: oracle build of the dictionary, its own text generator."""

    def __init__(self, name, keys=None):
        from oracle import oracle as O
        from aha_b200 import synth
        self.O, self.name, self.w = O, name, WORKLOADS[name]
        self.D = synth.doc_bytes(self.w["doc"])
        self.corpus, self.seed = CORPUS_ID[self.w["corpus"]], SEED_TEXT + self.w["seed"]
        keys = keys if keys is not None else make_patterns(name)
        if name == "c4":
            self.trie = O.Cedar()
            for k in keys:
                self.trie.insert(k)
            self.ac = None
        else:
            self.ac = O.AC.compile(keys)

    def text(self, n_docs):
        return self.O.gen_text(self.corpus, self.seed, 0, n_docs * self.D), np.arange(n_docs + 1, dtype=np.int64) * self.D

    def count(self, text, off, threads):
        if self.name == "c4":
            return self.trie.match_longest_count_mt(text, off, True, self.w["sep"], threads)
        if self.w["mode"] or self.w["sep"]:
            return self.ac.scan_mode_mt(text, off, threads, self.w["mode"], self.w["sep"])[0]
        return self.ac.scan_count_mt(text, off, threads)

    def stream_hash(self, text, off, threads):
        if self.name == "c4":
            return self.trie.match_longest_hash_mt(text, off, True, self.w["sep"], threads)
        if self.w["mode"] or self.w["sep"]:
            return self.ac.scan_mode_mt(text, off, threads, self.w["mode"], self.w["sep"])
        return self.ac.scan_hash_mt(text, off, threads)

    def timed(self, threads, budget_s, n_docs=None):
        """Counting walk over a bounded sample of the same stream, grown towards the time budget (at most 1 GiB)."""
        cap = max(1, (1 << 30) // self.D)
        n_docs = min(cap, n_docs or max(8, (32 << 20) // self.D))
        while True:
            text, off = self.text(n_docs)
            t0 = time.perf_counter()
            hits = self.count(text, off, threads)
            dt = time.perf_counter() - t0
            if dt >= budget_s / 4 or n_docs >= cap:
                break
            n_docs = min(cap, n_docs * 4)
        return {"gbs": n_docs * self.D / dt / 1e9, "hits_per_s": hits / dt, "bytes": n_docs * self.D, "docs": n_docs,
                "seconds": dt, "hits": hits}


def run_reference(args):
    """--impl reference: the oracle port of the reference's CPU path on all host threads.  Generates its own text
    (oracle/orc_gen_text) and never loads libaha_b200.so."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    from aha_b200 import synth
    synth.set_host_generator(lambda corpus, seed, first, n: O.gen_text(corpus, seed, first, n))
    threads = os.cpu_count() or 1
    cpu = CpuSide(args.workload)
    for _ in range(1 if args.warmup > 0 else 0):
        cpu.timed(threads, 2.0, 8)
    per, steps = None, []
    for _ in range(max(1, args.steps)):
        per = cpu.timed(threads, 20.0 / max(1, args.steps), per["docs"] if per else None)
        steps.append(per)
    gbs = float(np.mean([s["gbs"] for s in steps]))
    ms = float(np.mean([s["seconds"] for s in steps])) * 1e3
    w = WORKLOADS[args.workload]
    sample = "%d docs (%.0f MB) of the %s stream per step, %d threads, document-parallel, counting walk" % (
        steps[-1]["docs"], steps[-1]["bytes"] / 1e6, w["corpus"], threads)
    line = {
        "impl": "reference", "metric": "ac_scan_text_throughput", "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": w["desc"], "note": "CPU oracle port of the reference path; literal Crystal binary unavailable on this box"},
        "hits_per_s": float(np.mean([s["hits_per_s"] for s in steps])),
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def gpu_stream_hash(hits, doc_hit_off, n_docs):
    """Checksum of the first n_docs documents' hit streams, computed on the device (same formula as mix_hit)."""
    import torch
    off = doc_hit_off[:n_docs + 1]
    n = int(off[-1].item())
    if n == 0:
        return 0, 0
    total = 0
    step = 1 << 24
    for a in range(0, n, step):
        b = min(n, a + step)
        h = hits[a:b].to(torch.int64)
        idx = torch.arange(a, b, dtype=torch.int64, device=h.device)
        doc = torch.searchsorted(off, idx, right=True) - 1
        j = idx - off[doc]
        x = h[:, 0] * _KA + h[:, 1] * _KB + h[:, 2] * _KC + j * _KD + doc * _KE
        x = (x ^ (x >> 29)) * _KF
        total = (total + int(x.sum().item())) & _M64
    return n, total


def stats(xs):
    xs = [float(x) for x in xs]
    return {"min": min(xs), "median": float(np.median(xs)), "max": max(xs)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--bytes", type=int, default=0, help="override bytes per GPU (testing)")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--mode", default=None, choices=["match", "sep", "longest", "longest-intersectable"],
                    help="another entry point of the same dictionary and text (AC workloads): ACX#match(seq, sep=[' ']), "
                         "ACX#match_longest, match_longest(seq, true); default = the workload's own")
    ap.add_argument("--gather", action="store_true", help="also concatenate the ranks' hit streams on rank 0 (shard.gather_hits)")
    args = ap.parse_args()
    if args.mode and args.workload != "c4":
        m = {"match": (0, None), "sep": (0, [" "]), "longest": (LONGEST, None),
             "longest-intersectable": (LONGEST | INTERSECTABLE, None)}[args.mode]
        w0 = WORKLOADS[args.workload]
        WORKLOADS[args.workload] = dict(w0, mode=m[0], sep=m[1], desc=w0["desc"] + " [" + args.mode + "]")
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import aha_b200
    from aha_b200 import _lib, synth

    # stdout carries exactly one JSON line: everything else that native libraries write to fd 1 (NCCL prints its
    # version banner there) goes to stderr; the line itself is written to the saved descriptor at the end.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W = max(3, args.warmup)
    name = args.workload
    w = WORKLOADS[name]
    nbytes = args.bytes or w["nbytes"]
    D = synth.doc_bytes(w["doc"])
    n_docs = max(1, nbytes // D)
    total = n_docs * D
    corpus, seed = CORPUS_ID[w["corpus"]], SEED_TEXT + w["seed"]
    mode, sep = w["mode"], w["sep"]

    # ---- dictionary: built on rank 0 (host builder + flatten), broadcast as a blob over NCCL ----
    t0 = time.perf_counter()
    keys = None
    if rank == 0:
        keys = make_patterns(name)
        if name == "c4":
            trie = aha_b200.Cedar()
            for k in keys:
                trie.insert(k)
            ac = aha_b200.AC.compile(trie)
        else:
            ac = aha_b200.AC.compile(keys)
        blob = ac.blob_view() if world > 1 else None
    build_s = time.perf_counter() - t0
    bcast_ms = 0.0
    if world > 1:
        from aha_b200 import shard
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        buf = shard.broadcast_blob(blob if rank == 0 else b"", 0, "cuda")  # ncclBroadcast over NVLink
        e1.record()
        torch.cuda.synchronize()
        bcast_ms = e0.elapsed_time(e1)
        if rank != 0:
            ac = aha_b200.AC.from_device_blob(buf)
        del buf
    info = ac.info()

    # ---- corpus shard of this rank: documents [rank*n_docs, (rank+1)*n_docs) of the stream ----
    first = rank * total
    text = synth.gen_device(corpus, seed, first, total)
    doc_off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        return ac.scan_device(text, doc_off, mode, sep)

    # one result in flight: a caller releases a result (aha_b200_hits_free) before asking for the next
    res = None
    for _ in range(W):
        res = None
        res = step()
    n_hits = res.n_hits
    launches0 = _lib.lib().aha_b200_kernel_launches()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    scan_ms, walk_ms, expand_ms, wall_ms = [], [], [], []
    ev0.record()
    for _ in range(args.steps):
        res = None
        t_a = time.perf_counter()
        res = step()
        wall_ms.append((time.perf_counter() - t_a) * 1e3)
        scan_ms.append(res.scan_ms)
        walk_ms.append(res.walk_ms)
        expand_ms.append(res.expand_ms)
    ev1.record()
    barrier()
    gpu_launches = int(_lib.lib().aha_b200_kernel_launches() - launches0)  # kernels of libaha_b200.so in the timed region
    clocks = sampler.stop() if rank == 0 else None
    ms_total = ev0.elapsed_time(ev1)
    mine = [ms_total, float(n_hits)] + scan_ms + walk_ms + expand_ms + wall_ms
    t = torch.tensor(mine, dtype=torch.float64, device="cuda")
    if world > 1:
        tl = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(tl, t)  # per-GPU hit counts -> offsets of the concatenated result (document order)
        rows = [x.cpu().tolist() for x in tl]
    else:
        rows = [mine]
    K = args.steps
    per_rank = []
    for r, row in enumerate(rows):
        s_ms, w_ms, x_ms, wl_ms = row[2:2 + K], row[2 + K:2 + 2 * K], row[2 + 2 * K:2 + 3 * K], row[2 + 3 * K:2 + 4 * K]
        per_rank.append({"rank": r, "total_ms": row[0], "hits": int(row[1]), "scan_ms": stats(s_ms), "walk_ms": stats(w_ms),
                         "expand_ms": stats(x_ms), "call_wall_ms": stats(wl_ms),
                         "host_gap_ms": stats([a - b for a, b in zip(wl_ms, s_ms)])})
    ms_total = max(r["total_ms"] for r in per_rank)
    slowest = max(per_rank, key=lambda r: r["total_ms"])["rank"]
    hits_all = float(sum(r["hits"] for r in per_rank))
    ms_step = ms_total / args.steps
    value = world * total / (ms_step * 1e-3) / 1e9
    kms, wms, xms = float(np.mean(scan_ms)), float(np.mean(walk_ms)), float(np.mean(expand_ms))
    peak, peak_src = peaks()
    achieved = (total + 12.0 * n_hits) / (kms * 1e-3) / 1e9
    tr = measured_traffic(name)
    traffic = int(tr["dram_bytes_per_text_byte"] * total) if tr else None

    # ---- parity of the timed configuration: the first documents of rank 0's shard against the CPU leg ----
    parity = None
    cpu_side = None
    if rank == 0 and not args.no_cpu:
        cpu_side = CpuSide(name, keys)
        p_docs = min(n_docs, max(1, (64 << 20) // D))
        c_text, c_off = cpu_side.text(p_docs)
        hc, hash_cpu = cpu_side.stream_hash(c_text, c_off, os.cpu_count() or 1)
        hg, hash_gpu = gpu_stream_hash(res.hits_tensor(), res.doc_hit_off_tensor(), p_docs)
        parity = {"docs": p_docs, "bytes": p_docs * D, "hits_gpu": hg, "hits_cpu": int(hc), "hash_gpu": "%016x" % hash_gpu,
                  "hash_cpu": "%016x" % hash_cpu, "equal": bool(hg == hc and hash_gpu == hash_cpu)}

    # ---- ranks' streams concatenated on rank 0 in document order (optional: costs a D2H of every hit) ----
    gathered = None
    if args.gather and world > 1:
        from aha_b200 import shard
        g_docs = min(n_docs, 64)
        sub = ac.scan_device(text[:g_docs * D], doc_off[:g_docs + 1], mode, sep)
        h_loc = sub.hits_tensor().cpu().numpy().view(aha_b200.HIT_DTYPE).reshape(-1)
        o_loc = sub.doc_hit_off()
        hs, offs = shard.gather_hits(h_loc, o_loc, 0, "cuda")
        if rank == 0:
            gathered = {"docs": int(len(offs) - 1), "hits": int(len(hs)), "monotone": bool(np.all(np.diff(offs) >= 0)),
                        "first_block_equal": bool(hs[:len(h_loc)].tobytes() == h_loc.tobytes())}
        sub = None
    res = None

    # ---- end to end through the C ABI with host buffers (rank-local shard, pinned host text) ----
    if args.e2e_steps <= 0:
        args.e2e_steps, e2e_skip = 1, True
    else:
        e2e_skip = False
    e2e_bytes = min(total, (1 << 30) // D * D) if not args.bytes else total
    if e2e_skip:
        e2e_bytes = D
    e2e_docs = e2e_bytes // D
    host_text = torch.empty(e2e_bytes, dtype=torch.uint8, pin_memory=True)
    host_text.copy_(text[:e2e_bytes])
    host_np = host_text.numpy()
    host_off = np.arange(e2e_docs + 1, dtype=np.int64) * D
    for _ in range(2):  # warm-up: pinned result buffers and the device buffers reach their steady size
        r = ac.scan_batch((host_np, host_off), mode, sep)
        e2e_hits = r.n_hits
        r = None  # a caller releases one result before asking for the next (aha_b200_hits_free)
    barrier()
    e2e_ms = []
    for _ in range(args.e2e_steps):
        t_a = time.perf_counter()
        r = ac.scan_batch((host_np, host_off), mode, sep)
        e2e_hits = r.n_hits  # hits and doc_hit_off are in pinned host memory when the call returns
        r = None
        e2e_ms.append((time.perf_counter() - t_a) * 1e3)
    torch.cuda.synchronize()
    e2e_s = float(np.mean(e2e_ms)) * 1e-3
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e = {"value": world * e2e_bytes / float(e2e_t.item()) / 1e9, "unit": "GB/s",
           "h2d_bytes_per_step": int(e2e_bytes + 8 * (e2e_docs + 1)),
           "d2h_bytes_per_step": int(12 * e2e_hits + 8 * (e2e_docs + 1)),
           "call_ms": stats(e2e_ms),
           "note": "aha_b200_scan (C ABI, pinned host text in, pinned host hits out, H2D/scan/D2H pipelined over 64 MB runs) "
                   "on %d MB of host text per rank per step" % (e2e_bytes >> 20)}

    if rank == 0:
        cpu = None
        if cpu_side is not None:
            threads = os.cpu_count() or 1
            c = cpu_side.timed(threads, 15.0)
            c1 = cpu_side.timed(1, 6.0, 8)
            cpu = {"value": c["gbs"], "unit": "GB/s", "cores": threads, "kind": "port",
                   "sample": "%d docs (%.0f MB) of the same stream, %d threads document-parallel, counting walk; 1 thread: %.3f GB/s" % (
                       c["docs"], c["bytes"] / 1e6, threads, c1["gbs"]),
                   "hits_per_s": c["hits_per_s"]}
        dominant = None
        if wms > 0:
            dominant = {"name": "char_walk_kernel", "ms": wms, "share_of_step": wms / kms,
                        "second": {"name": "char_expand_kernel", "ms": xms, "share_of_step": xms / kms}}
        line = {
            "metric": "ac_scan_text_throughput", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": w["desc"], "bytes_per_gpu": total, "docs_per_gpu": n_docs, "n_patterns": info["n_keys"],
                       "dfa_states": info["n_states"], "byte_classes": info["n_classes"], "hot_states": info["hot_states"],
                       "max_key_len": info["max_key_len"], "hits_per_gpu_step": int(n_hits), "mode": mode | (SEP if sep else 0),
                       "l2": "input (%.1f GB) larger than L2, no flush needed" % (total / 1e9),
                       "results_in_flight": 1,
                       "impl": "char walk -> scan -> expand" if wms > 0 else ("per-document machines" if mode & (LONGEST | TRIE_LONGEST) else "fused byte DFA")},
            "hits_per_s": hits_all / (ms_step * 1e-3),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": (tr or {}).get("source"), "peak_source": peak_src, "kernel_ms": kms,
                         "algorithmic_bytes": "1 B read per text byte + 12 B written per hit",
                         "kernel": "all scan launches of one step (walk, prefix scan, expand, finalize), CUDA events in the library",
                         "dominant_kernel": dominant},
            "per_rank": per_rank, "slowest_rank": slowest,
            "parity": parity, "gathered": gathered,
            "cpu_baseline": cpu,
            "e2e": e2e, "gpu_launches": gpu_launches, "clocks": clocks,
            "dfa_build_s": build_s, "dfa_broadcast_ms": bcast_ms,
        }
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
