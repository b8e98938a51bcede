#!/usr/bin/env python
"""Per-instruction view of an .ncu-rep (first captured launch): shared-memory wavefronts per opcode
and the instructions with the most long/short-scoreboard stall samples.
    python scripts/ncu_source.py gpurun_out/<tag>/prof.ncu-rep"""
import csv, collections, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
R = []
for r in rows[2:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        if R: break
        continue
    R.append(r)
def col(name): return hdr.index(name)
def num(r, i):
    try: return int(r[i] or 0)
    except ValueError: return 0
i_src, iw, ii, iex = col("Source"), col("L1 Wavefronts Shared"), col("L1 Wavefronts Shared Ideal"), col("Instructions Executed")
print("instructions:", len(R), " executed warp-instr:", sum(num(r, iex) for r in R))
agg = collections.defaultdict(lambda: [0, 0, 0])
for r in R:
    w = num(r, iw)
    if not w: continue
    s = r[i_src].strip().split()
    o = s[1] if s[0].startswith("@") else s[0]
    a = agg[o]; a[0] += w; a[1] += num(r, ii); a[2] += num(r, iex)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"  {k:10s} wavefronts {v[0]:9d} ideal {v[1]:9d} executed {v[2]:8d} wf/inst {v[0]/max(v[2],1):.2f}")
tot = sum(num(r, col("Warp Stall Sampling (All Samples)")) for r in R)
print("stall samples:", tot)
for name in ("stall_long_sb", "stall_short_sb", "stall_barrier", "stall_wait", "stall_mio", "stall_lg", "stall_branch_resolving", "stall_selected", "stall_not_selected", "stall_math", "stall_membar"):
    c = col(name)
    d = sorted(((num(r, c), k, r[i_src].strip()) for k, r in enumerate(R)), reverse=True)
    print(f"{name}: {sum(x[0] for x in d)}")
    if name in ("stall_long_sb", "stall_short_sb", "stall_barrier"):
        for s, k, src in d[:6]: print(f"     {s:5d} #{k:5d} {src[:72]}")
