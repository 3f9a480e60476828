#!/usr/bin/env python
"""A/B of library builds on one GPU box.

    python tools/ab.py build tag1:-DFLAG_A tag2:-DFLAG_A,-DFLAG_B ...   (here: cross-compiles aha_b200/libaha_b200_<tag>.so)
    python tools/ab.py run [--workload c2] tag1 tag2 ...               (on the box: short bench per build, one line each)
"tag:" alone is the default build.  Every build runs the same bench command in its own process."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def so(tag):
    return os.path.join(ROOT, "aha_b200", "libaha_b200_%s.so" % tag)


def main():
    cmd, args = sys.argv[1], sys.argv[2:]
    if cmd == "build":
        from aha_b200 import _lib
        procs = []
        for a in args:
            tag, _, flags = a.partition(":")
            flags = [f for f in flags.split(",") if f]
            print("build", tag, flags, flush=True)
            _lib.build(out=so(tag), extra_flags=flags)
        return
    workload, reps, nbytes, steps, mode = "c2", 2, "0", "20", "match"
    while args and args[0].startswith("--"):
        if args[0] == "--workload":
            workload = args[1]
        elif args[0] == "--reps":
            reps = int(args[1])
        elif args[0] == "--bytes":
            nbytes = args[1]
        elif args[0] == "--steps":
            steps = args[1]
        elif args[0] == "--mode":
            mode = args[1]
        args = args[2:]
    for rep in range(reps):
        for tag in args:
            env = dict(os.environ, AHA_B200_SO=so(tag))
            out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", workload, "--bytes", nbytes, "--steps", steps, "--warmup", "3",
                                  "--no-cpu", "--e2e-steps", "0", "--mode", mode], env=env, capture_output=True, text=True)
            try:
                d = json.loads(out.stdout.strip().splitlines()[-1])
                r = d["per_rank"][0]
                med = lambda k: (r.get(k) or {}).get("median") or 0.0
                print("%-12s %7.1f GB/s  step %.3f  scan %.3f  walk %.3f  expand %.3f  gap %.3f  hits %d" % (
                    tag, d["value"], d["ms_per_step"], med("scan_ms"), med("walk_ms"), med("expand_ms"),
                    med("host_gap_ms"), d["config"]["hits_per_gpu_step"]), flush=True)
            except Exception:
                print(tag, "FAILED", out.stderr[-600:], flush=True)


if __name__ == "__main__":
    main()
