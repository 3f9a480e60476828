"""Static look at the hot loop of char_walk_kernel: distances between the dp4a decode groups in the SASS."""
import subprocess, sys, re
fn = sys.argv[1] if len(sys.argv) > 1 else "_ZN3aha16char_walk_kernelENS_11SplitParamsE"
out = subprocess.run(["cuobjdump", "-sass", "-fun", fn, "aha_b200/libaha_b200.so"], capture_output=True, text=True).stdout
ins = [l for l in out.splitlines() if re.search(r"/\*[0-9a-f]{4,}\*/", l)]
ins = [re.sub(r"/\*[0-9a-f]+\*/", "", l).strip().rstrip(";").strip() for l in ins]
ins = [i for i in ins if i]
print("instructions:", len(ins))
idx = [i for i, l in enumerate(ins) if "IDP.4A" in l]
print("dp4a at:", idx)
if "-p" in sys.argv and idx:
    a = int(sys.argv[sys.argv.index("-p") + 1]); b = int(sys.argv[sys.argv.index("-p") + 2])
    for i in range(a, b): print(i, ins[i])
