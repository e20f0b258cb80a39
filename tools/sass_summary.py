"""Instruction-mnemonic summary of the kernels in libfluidsim_b200.so (cuobjdump -sass), written to
profiles/<tag>_sass_summary.md: which memory / synchronisation instructions each kernel really contains.
usage: python tools/sass_summary.py r02"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
lib = os.path.join(ROOT, "fluid_sim_b200", "libfluidsim_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
WATCH = ["LDGSTS", "LDGDEPBAR", "DEPBAR", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "LDS", "STS", "LDG", "STG", "LDL", "STL", "SHFL",
         "BAR", "MEMBAR", "ERRBAR", "CCTL", "ATOMG", "ATOMS", "RED", "FADD", "FMUL", "FFMA", "MUFU", "NANOSLEEP", "BPT", "CS2R", "S2UR"]
rows = []
for part in re.split(r"\n\s*Function : ", sass)[1:]:
    name = part.split("\n", 1)[0].strip()
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    dem = re.sub(r"fsb::\(anonymous namespace\)::", "", dem)
    dem = re.sub(r"\(.*", "", dem)
    ops = collections.Counter()
    total = 0
    for line in part.splitlines():
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if not m:
            continue
        total += 1
        ops[m.group(1).split(".")[0]] += 1
    rows.append((dem, total, ops))
rows.sort(key=lambda r: -r[1])
out = ["# SASS summary of `libfluidsim_b200.so` (sm_100a, `cuobjdump -sass`; static instruction counts)", "",
       "Built with `-fmad=false`: every FFMA below belongs to a division / square-root / conversion sequence the compiler",
       "expands (`__fdiv_rn`, `sqrtf`, integer division), not to a contracted multiply-add of the reference's arithmetic.",
       "Staging into shared memory is `cp.async` (LDGSTS); there is no TMA instruction (UTMALDG / UBLKCP): one bulk copy",
       "per 288-byte row was measured 2.5x slower than plain loads in the pointwise phases (DESIGN.md section 4).", "",
       "| kernel | instructions | " + " | ".join(WATCH) + " |", "|---|---|" + "---|" * len(WATCH)]
for dem, total, ops in rows:
    if total < 50:
        continue
    out.append(f"| `{dem[:70]}` | {total} | " + " | ".join(str(ops.get(w, 0)) for w in WATCH) + " |")
open(os.path.join(ROOT, "profiles", f"{tag}_sass_summary.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out[:14]))
