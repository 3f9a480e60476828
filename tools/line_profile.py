"""Joins an ncu SASS source page (csv) with nvdisasm line info: executed instructions and stall samples per source line.

usage: line_profile.py <ncu source csv> <kernel mangled name> [top N]
The cubin is taken from aha_b200/libaha_b200.so (must be the build that was profiled)."""
import csv, collections, os, re, subprocess, sys, tempfile
src_csv, fn = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath("aha_b200/libaha_b200.so")], cwd=d, capture_output=True)
txt = []
for cub in sorted(f for f in os.listdir(d) if f.endswith(".cubin")):  # one cubin per translation unit
    t = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, cub)], capture_output=True, text=True).stdout.splitlines()
    if any(l.startswith(".text." + fn + ":") for l in t):
        txt = t
# find the function's text section
start = max(i for i, l in enumerate(txt) if l.startswith(".text." + fn + ":"))
lines, cur = [], None
for l in txt[start + 1:]:
    if l.startswith("\t.section") or l.startswith(".text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.search(r"/\*[0-9a-f]{4,}\*/", l) and not l.strip().startswith(".") and ";" in l:
        lines.append(cur)
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
data = [r for r in rows[2:] if len(r) > 10 and r[0].startswith("0x")]
iE, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
print("sass instr (ncu) %d, (nvdisasm) %d" % (len(data), len(lines)))
n = min(len(data), len(lines))
ex, sm = collections.Counter(), collections.Counter()
for i in range(n):
    ex[lines[i]] += int(data[i][iE]); sm[lines[i]] += int(data[i][iS])
te, ts = sum(ex.values()), sum(sm.values())
print("total executed %d  samples %d" % (te, ts))
for k, v in ex.most_common(top):
    print("%-22s %5s  exec %5.2f%%  samples %5.2f%%" % (k[0] if k else "?", k[1] if k else "", 100.0 * v / te, 100.0 * sm[k] / ts))
