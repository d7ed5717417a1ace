"""Static SASS instruction mix per kernel of libgmpm_b200.so -> profiles/rNN_sass_summary.txt.
  python tools/sass_summary.py [out]"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "drake_b200", "lib", "libgmpm_b200.so")
out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_summary.txt")
txt = subprocess.check_output(["cuobjdump", "-sass", lib]).decode()
funcs = re.split(r"\n\s*Function : ", txt)[1:]
pats = ["LDGSTS", "UBLKCP", "UTMALDG", "UTMASTG", "UTCHMMA", "UTCQMMA", "ATOMG", "ATOMS", r"RED\.", r"LDS\.128", r"LDS\.64",
        "LDS ", r"STS\.128", r"LDG\.E\.128", r"LDG\.E\.64", r"STG\.E\.128", r"STG\.E\.64", "DFMA", "DMUL", "DADD", "FFMA",
        "SHFL", "MUFU", r"BAR\.SYNC", "CCTL", "MEMBAR", "FENCE"]
rows = []
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    short = re.sub(r"\(.*", "", dem).replace("void ", "").replace("gmpm::", "")
    n_inst = len(re.findall(r"/\*[0-9a-f]{4}\*/", f))
    rows.append((short, n_inst, {p: len(re.findall(r"\b" + p, f)) for p in pats}))
rows.sort()
out = ["# SASS summary of drake_b200/lib/libgmpm_b200.so (cuobjdump -sass, sm_100a, nvcc 12.9; static instruction counts per kernel)",
       "# Blackwell-specific: UBLKCP = cp.async.bulk (bulk copy engine; p2g_col_kernel<..,true>), LDGSTS = cp.async (G2P tile staging).",
       "# No UTMALDG/UTMASTG (tensor-map TMA) and no UTC*MMA (tcgen05): the path has no GEMM-shaped operand (north_star).",
       "%-58s %6s " % ("kernel", "instr") + " ".join("%7s" % p.replace("\\", "").strip() for p in pats)]
for short, n, c in rows:
    out.append("%-58s %6d " % (short[:58], n) + " ".join("%7d" % c[p] for p in pats))
open(out_path, "w").write("\n".join(out) + "\n")
print(out_path, len(rows), "kernels")
