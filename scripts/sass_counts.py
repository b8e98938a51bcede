#!/usr/bin/env python
"""Per-kernel SASS instruction counts of libfluid_cuda.so (cuobjdump -sass): total, TMA (UTMALDG), mbarrier
(SYNCS), FFMA (must be 0 on the hot path: --fmad=false is the parity contract), and where the FFMAs that exist
live.  Writes profiles/r02_sass_tma.txt."""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "fluid_simulation_b200", "libfluid_cuda.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
per, name = {}, None
keys = ("UTMALDG", "SYNCS", "FFMA", "LDS", "STS", "SHFL", "BAR")
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1); per[name] = dict.fromkeys(keys, 0); per[name]["n"] = 0
    elif name and re.match(r"\s+/\*[0-9a-f]{4,6}\*/", line):
        per[name]["n"] += 1
        for k in keys:
            if re.search(r"\b" + k + r"\b", line):
                per[name][k] += 1
def demangle(n):
    try:
        return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
    except Exception:
        return n
lines = ["# cuobjdump -sass fluid_simulation_b200/libfluid_cuda.so (sm_100a only), instructions per kernel",
         "# %-78s %7s %8s %6s %5s %5s %5s %5s" % ("kernel", "instr", "UTMALDG", "SYNCS", "FFMA", "LDS", "STS", "SHFL")]
tot = dict.fromkeys(keys, 0)
for n in sorted(per, key=lambda k: -per[k]["n"]):
    d = per[n]
    for k in keys: tot[k] += d[k]
    lines.append("%-80s %7d %8d %6d %5d %5d %5d %5d" % (demangle(n)[:80], d["n"], d["UTMALDG"], d["SYNCS"], d["FFMA"], d["LDS"], d["STS"], d["SHFL"]))
lines.append("# totals: " + ", ".join(f"{k} {tot[k]}" for k in keys))
ff = [demangle(n) for n in per if per[n]["FFMA"]]
lines.append("# kernels that contain FFMA (IEEE division / cbrt / double sequences and index arithmetic off the per-step path): "
             + ("; ".join(ff) if ff else "none"))
open(os.path.join(ROOT, "profiles", "r02_sass_tma.txt"), "w").write("\n".join(lines) + "\n")
print("\n".join(lines[-3:]))
