#!/usr/bin/env python
"""per-kernel SASS mnemonic counts of libaha_b200.so (cuobjdump -sass): the static evidence for TMA bulk copies (UBLKCP),
mbarrier waits (SYNCS), shared-memory lookups (LDS) and global traffic (LDG / STG) in each kernel.
usage: tools/sass_table.py [out.md]"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "aha_b200", "libaha_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
dem = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
cols = ["UBLKCP", "SYNCS", "LDS", "STS", "LDG", "STG", "LDC", "SHFL", "BAR", "ATOM", "RED", "BRA"]
rows, cur, cnt = [], None, None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if cur: rows.append((cur, cnt))
        cur, cnt = m.group(1), collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        cnt["total"] += 1
        for c in cols:
            if op.split(".")[0] == c or (c == "ATOM" and op.startswith("ATOM")):
                cnt[c] += 1
if cur: rows.append((cur, cnt))
out = ["| kernel | SASS instr | " + " | ".join(cols) + " |", "|---|---:|" + "---:|" * len(cols)]
for name, c in sorted(rows, key=lambda r: -r[1]["total"]):
    d = dem(name)
    d = re.sub(r"\((?:aha::)?(?:Split|Fused|Scan|Na|int|const|unsigned).*", "", d).replace("aha::", "").replace("void ", "")
    if d.startswith("cub::") or "thrust" in d: continue
    out.append("| `%s` | %d | %s |" % (d[:70], c["total"], " | ".join(str(c[k]) for k in cols)))
text = "\n".join(out) + "\n"
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write("# SASS mnemonic counts per kernel (sm_100a, `cuobjdump -sass aha_b200/libaha_b200.so`)\n\n"
                                 "UBLKCP = `cp.async.bulk` (TMA bulk copy global -> shared), SYNCS = mbarrier arrive/try_wait, LDS/STS = shared-memory "
                                 "loads/stores, LDG/STG = global loads/stores.  Static counts of the compiled code, not executed counts.\n\n" + text)
print(text)
