"""Copies the artefacts of tools/final_run.sh from gpurun_out/ into profiles/ and rewrites profiles/traffic.json and a
summary (stdout) from the ncu capture.  usage: update_profiles.py [tag]"""
import csv, json, os, shutil, subprocess, sys
tag = sys.argv[1] if len(sys.argv) > 1 else "r01_final"
for f in ("bench.json", "reference.json", "launches.csv"):
    shutil.copy("gpurun_out/%s_%s" % (tag, f), "profiles/%s_%s" % (tag, f))
raw = subprocess.run(["ncu", "-i", "gpurun_out/%s_walk_expand_full.ncu-rep" % tag, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
open("profiles/%s_walk_expand_ncu_full_raw.csv" % tag, "w").write(raw)
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
bench = json.loads(open("profiles/%s_bench.json" % tag).read())
total = bench["config"]["bytes_per_gpu"]
hits = bench["config"]["hits_per_gpu_step"]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct"]
per = {}
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")].split("(")[0]
    d = {}
    for w in want:
        if w not in hdr:
            continue
        i = hdr.index(w)
        d[w] = (float(r[i]), units[i])
    per[name] = d
    print(name)
    for w in want:
        if w in d:
            print("   %-70s %14.3f %s" % (w, d[w][0], d[w][1]))
def gb(d, k):
    v, u = d[k]
    return v * {"Gbyte": 1.0, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9}[u]
tr = {"c2": {"dram_bytes_per_text_byte": sum(gb(d, "dram__bytes_read.sum") + gb(d, "dram__bytes_write.sum") for d in per.values()) * 1e9 / total,
             "source": "profiles/%s_walk_expand_ncu_full_raw.csv (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, char_walk_kernel + "
                       "char_expand_kernel of one C2 step; the other launches of a step move < 0.02 GB)" % tag,
             "per_kernel_gb": {k: {"read": gb(d, "dram__bytes_read.sum"), "write": gb(d, "dram__bytes_write.sum"),
                                   "ms": d["gpu__time_duration.sum"][0]} for k, d in per.items()},
             "algorithmic_gb": (total + 12 * hits) / 1e9}}
json.dump(tr, open("profiles/traffic.json", "w"), indent=1)
print(json.dumps(tr, indent=1))
