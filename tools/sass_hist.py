"""Per-kernel SASS opcode histogram of libbt_b200.so -> profiles/r2_sass_opcodes.txt (needs cuobjdump and c++filt; no GPU)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "bayesian-torch_b200", "csrc", "libbt_b200.so")
out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r2_sass_opcodes.txt")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pat = re.compile(r"^\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)")
fn, hist = None, collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    if "Function :" in line:
        fn = line.split(":", 1)[1].strip()
        continue
    m = pat.match(line)
    if m and fn:
        hist[fn][m.group(1).split(".")[0]] += 1
WANT = ["UTCHMMA", "UTCQMMA", "UTCMMA", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "LDTM", "STTM", "UTCBAR", "MUFU", "HMMA", "FFMA",
        "ACQBULK", "CCTL", "UTMACCTL", "UTMACMDFLUSH", "FENCE", "ELECT", "USETMAXREG", "BAR", "ATOM", "RED", "LDG", "STG", "LDS", "STS"]
names = list(hist)
dem = subprocess.run(["c++filt", *names], capture_output=True, text=True).stdout.splitlines()
out = ["SASS opcode histogram of libbt_b200.so (cuobjdump -sass, sm_100a), per kernel: total instructions and the opcode families that show which",
       "hardware path a kernel uses (UTCHMMA = tcgen05.mma, UTMALDG/UTMASTG = TMA load/store, LDTM = tcgen05.ld, SYNCS = mbarrier, UBLKCP = bulk copy (DSMEM),",
       "MUFU = special-function unit (Box-Muller / softplus), ACQBULK/USETMAXREG = PDL / setmaxnreg).  Regenerate: python tools/sass_hist.py", ""]
for n, d in zip(names, dem):
    d = re.sub(r"\(anonymous namespace\)::", "", d).split("(")[0][:110]
    out.append(f"{sum(hist[n].values()):7d}  {d}")
    out.append("         " + "  ".join(f"{k}={hist[n][k]}" for k in WANT if hist[n].get(k)))
open(out_path, "w").write("\n".join(out) + "\n")
print(len(names), "kernels ->", out_path)
