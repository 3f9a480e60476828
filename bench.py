#!/usr/bin/env python
"""bench.py -- AC scan GB/s of text (+ hits/s) on N x B200, beside Aha::AC on the host cores.

Workload (BASELINE.json configs[1], "C2"): 10k CJK UTF-8 patterns over 2 GB of synthetic Chinese text
per GPU (2048 documents of ~1 MiB), dictionary sampled from the same stream.  One step = one pass of
the scan over the whole batch.  `value` times the scan with text resident in HBM; `e2e` times the
C-ABI call aha_b200_scan() on HOST buffers (H2D text, scan, D2H hits inside the timed region).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c1|c2|c3|c5]
N > 1 is launched by torchrun (one rank per GPU): the corpus shards by document (weak scaling), the DFA
is built on rank 0 and broadcast over NCCL, hit counts are all-gathered.
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

WORKLOADS = {
    # name: (corpus, n_patterns, bytes per GPU, doc bytes, description)
    "c1": ("EN", 1000, 100 << 20, 1 << 20, "C1: 1k ASCII word patterns / 100 MB EN text"),
    "c2": ("ZH", 10000, 2 << 30, 1 << 20, "C2: 10k CJK patterns / 2 GB ZH text"),
    "c3": ("MIX", 100000, 10 << 30, 1 << 20, "C3: 100k patterns / 10 GB MIX text"),
    # configs[4]: 100 GB over 8 GPUs = 12.5 GB per GPU, the 3.2 GB byte DFA broadcast over NCCL
    "c5": ("MIX", 1000000, 12500 << 20, 1 << 20, "C5: 1M patterns / 100 GB MIX text document-sharded over 8 GPUs (12.5 GB per GPU)"),
}
CORPUS = {"c1": ("EN", 0), "c2": ("ZH", 1), "c3": ("MIX", 2), "c5": ("MIX", 2)}  # corpus, seed offset


def make_patterns(name):
    from aha_b200 import synth
    if name == "c1":
        return synth.patterns_en_words(1000), synth.EN, synth.SEED_TEXT
    if name == "c2":
        return synth.patterns_zh_ngrams(10000), synth.ZH, synth.SEED_TEXT + 1
    return synth.patterns_mix(WORKLOADS[name][1]), synth.MIX, synth.SEED_TEXT + 2


def measured_traffic(name):
    """dram__bytes_read.sum + dram__bytes_write.sum of the scan kernel from the committed `ncu --set full` capture
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


def cpu_leg(name, threads, budget_s, sample_docs=None):
    """Times the CPU oracle (restatement of ACX#match, ac.cr:176-192,265-286) on a bounded sample."""
    from aha_b200 import synth
    from oracle import oracle as O
    keys, corpus, seed = make_patterns(name)
    D = synth.doc_bytes(WORKLOADS[name][3])
    o = O.AC.compile(keys)
    n_docs = sample_docs or 32
    text = synth.gen_host(corpus, seed, 0, n_docs * D)
    off = np.arange(n_docs + 1, dtype=np.int64) * D
    t0 = time.perf_counter()
    hits = o.scan_count_mt(text, off, threads)
    dt = time.perf_counter() - t0
    # grow the sample towards the time budget (bounded: at most 1 GiB)
    while dt < budget_s / 4 and n_docs * D < (1 << 30):
        n_docs *= 4
        text = synth.gen_host(corpus, seed, 0, n_docs * D)
        off = np.arange(n_docs + 1, dtype=np.int64) * D
        t0 = time.perf_counter()
        hits = o.scan_count_mt(text, off, threads)
        dt = time.perf_counter() - t0
    return {"gbs": n_docs * D / dt / 1e9, "hits_per_s": hits / dt, "bytes": n_docs * D, "docs": n_docs, "seconds": dt,
            "hits": hits}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    steps = []
    for _ in range(args.warmup if args.warmup < 2 else 1):
        cpu_leg(args.workload, threads, 2.0, 8)
    per = None
    for _ in range(max(1, args.steps)):
        per = cpu_leg(args.workload, threads, 20.0 / max(1, args.steps), per["docs"] if per else None)
        steps.append(per)
    gbs = float(np.mean([s["gbs"] for s in steps]))
    ms = float(np.mean([s["seconds"] for s in steps])) * 1e3
    sample = "%d docs (%.0f MB) of the %s stream per step, %d threads, document-parallel" % (
        steps[-1]["docs"], steps[-1]["bytes"] / 1e6, WORKLOADS[args.workload][0], threads)
    line = {
        "impl": "reference", "metric": "ac_scan_text_throughput", "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][4], "note": "CPU oracle port of Aha::AC#match; literal Crystal binary unavailable on this box"},
        "hits_per_s": float(np.mean([s["hits_per_s"] for s in steps])),
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--bytes", type=int, default=0, help="override bytes per GPU (testing)")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import aha_b200
    from aha_b200 import synth

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
    corpus_name, n_pat, nbytes, doc_target, desc = WORKLOADS[name]
    if args.bytes:
        nbytes = args.bytes
    D = synth.doc_bytes(doc_target)
    n_docs = max(1, nbytes // D)
    total = n_docs * D

    # ---- dictionary: built on rank 0 (host builder + flatten), broadcast as a blob over NCCL ----
    t0 = time.perf_counter()
    if rank == 0:
        keys, corpus, seed = make_patterns(name)
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
            corpus, seed = getattr(synth, CORPUS[name][0]), synth.SEED_TEXT + CORPUS[name][1]
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
        return ac.scan_device(text, doc_off)

    # Results in flight.  With the previous result still alive during the next scan the hit buffers alternate between
    # two pool blocks (C2: 632-643 GB/s against 592-620 GB/s when the block just freed is taken again, same box, A/B in
    # tools/chunk_ab.sh).  With multi-GB results (C3, C5) two alive make the stream-ordered pool grow through the
    # first calls, which puts cudaMallocAsync's mapping of the buffers into the timed steps (tools/gap_probe.py), so
    # there the previous result is released first.  AHA_BENCH_KEEP_PREV=0|1 forces either.
    res = step()
    keep_prev = 12 * res.n_hits < (2 << 30)
    if os.environ.get("AHA_BENCH_KEEP_PREV") in ("0", "1"):
        keep_prev = os.environ["AHA_BENCH_KEEP_PREV"] == "1"
    for _ in range(W - 1):
        if not keep_prev:
            res = None
        res = step()
    n_hits = res.n_hits
    from aha_b200 import _lib
    launches0 = _lib.lib().aha_b200_kernel_launches()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms, walk_ms = [], []
    ev0.record()
    for _ in range(args.steps):
        if not keep_prev:
            res = None
        res = step()
        kernel_ms.append(res.scan_ms)
        walk_ms.append(res.walk_ms)
    ev1.record()
    barrier()
    gpu_launches = int(_lib.lib().aha_b200_kernel_launches() - launches0)  # kernels of libaha_b200.so in the timed region
    clocks = sampler.stop() if rank == 0 else None
    ms_total = ev0.elapsed_time(ev1)
    t = torch.tensor([ms_total, float(n_hits)], dtype=torch.float64, device="cuda")
    if world > 1:
        tl = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(tl, t)  # per-GPU hit counts -> offsets of the concatenated result (document order)
        ms_total = max(float(x[0]) for x in tl)
        hits_all = sum(float(x[1]) for x in tl)
    else:
        hits_all = float(n_hits)
    ms_step = ms_total / args.steps
    value = world * total / (ms_step * 1e-3) / 1e9
    kms = float(np.mean(kernel_ms))
    wms = float(np.mean(walk_ms))
    peak, peak_src = peaks()
    achieved = (total + 12.0 * n_hits) / (kms * 1e-3) / 1e9
    tr = measured_traffic(name)
    traffic = int(tr["dram_bytes_per_text_byte"] * total) if tr else None

    # ---- end to end through the C ABI with host buffers (rank-local shard, pinned host text) ----
    e2e = None
    e2e_bytes = min(total, (1 << 30) // D * D) if not args.bytes else total
    e2e_docs = e2e_bytes // D
    host_text = torch.empty(e2e_bytes, dtype=torch.uint8, pin_memory=True)
    host_text.copy_(text[:e2e_bytes])
    host_np = host_text.numpy()
    host_off = np.arange(e2e_docs + 1, dtype=np.int64) * D
    for _ in range(2):  # warm-up: pinned result buffers and the device pool reach their steady size
        r = ac.scan_batch((host_np, host_off))
        e2e_hits = r.n_hits
        r = None  # a caller releases one result before asking for the next (aha_b200_hits_free)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        r = ac.scan_batch((host_np, host_off))
        e2e_hits = r.n_hits  # hits and doc_hit_off are in pinned host memory when the call returns
        r = None
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / args.e2e_steps
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e = {"value": world * e2e_bytes / float(e2e_t.item()) / 1e9, "unit": "GB/s",
           "h2d_bytes_per_step": int(e2e_bytes + 8 * (e2e_docs + 1)),
           "d2h_bytes_per_step": int(12 * e2e_hits + 8 * (e2e_docs + 1)),
           "note": "aha_b200_scan (C ABI, pinned host text in, pinned host hits out, H2D/scan/D2H pipelined over 64 MB runs) "
                   "on %d MB of host text per rank per step" % (e2e_bytes >> 20)}

    if rank == 0:
        cpu = None
        if not args.no_cpu:
            threads = os.cpu_count() or 1
            c = cpu_leg(name, threads, 15.0)
            c1 = cpu_leg(name, 1, 6.0, 8)
            cpu = {"value": c["gbs"], "unit": "GB/s", "cores": threads, "kind": "port",
                   "sample": "%d docs (%.0f MB) of the same stream, %d threads document-parallel; 1 thread: %.3f GB/s" % (
                       c["docs"], c["bytes"] / 1e6, threads, c1["gbs"]),
                   "hits_per_s": c["hits_per_s"]}
        line = {
            "metric": "ac_scan_text_throughput", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": desc, "bytes_per_gpu": total, "docs_per_gpu": n_docs, "n_patterns": info["n_keys"],
                       "dfa_states": info["n_states"], "byte_classes": info["n_classes"], "hot_states": info["hot_states"],
                       "max_key_len": info["max_key_len"], "hits_per_gpu_step": int(n_hits),
                       "l2": "input (%.1f GB) larger than L2, no flush needed" % (total / 1e9),
                       "results_in_flight": 2 if keep_prev else 1,
                       "impl": "two-pass" if os.environ.get("AHA_B200_IMPL") == "0" else (
                           "char walk -> scan -> expand" if wms > 0 else "fused byte DFA")},
            "hits_per_s": hits_all / (ms_step * 1e-3),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": (tr or {}).get("source"), "peak_source": peak_src, "kernel_ms": kms,
                         "algorithmic_bytes": "1 B read per text byte + 12 B written per hit",
                         "kernel": "all scan launches of one step (walk, prefix scan, expand, finalize), CUDA events in the library",
                         "dominant_kernel": {"name": "char_walk_kernel", "ms": wms, "share_of_step": wms / kms} if wms > 0 else None},
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
