#!/usr/bin/env python
"""Turn gpurun_out/<tag> (launches.csv from the gpu__time_duration pass, prof*.ncu-rep from --set full)
into the tracked summaries under profiles/.
    python scripts/make_profiles.py <round-tag> <gpurun_out/dir with launches.csv> <ncu-rep> [<ncu-rep> ...]"""
import collections, csv, io, os, shutil, subprocess, sys

tag, ldir, reps = sys.argv[1], sys.argv[2], sys.argv[3:]
os.makedirs("profiles", exist_ok=True)
lc = os.path.join(ldir, "launches.csv")
if os.path.exists(lc):
    shutil.copy(lc, f"profiles/{tag}_launches.csv")
    rows = list(csv.reader(open(lc)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr = rows[hi]; kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[hi + 1:]:
        if len(r) <= mv: continue
        a = agg[r[kn].split("(StreamArgs")[0].split("(Grid")[0]]; a[0] += 1; a[1] += float(r[mv].replace(",", ""))
    tot = sum(v[1] for v in agg.values())
    with open(f"profiles/{tag}_launch_shares.txt", "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none, python bench.py --steps 1 --warmup 3 --no-cpu\n"
                f"# (cold-cache, serialised launches: compare SHARES, not absolutes); unit {rows[hi+1][mu]}\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k:64s} n={v[0]:4d} total={v[1]:13.1f} share={100*v[1]/tot:5.1f}% avg={v[1]/v[0]:10.1f}\n")
for rep in reps:
    name = os.path.basename(os.path.dirname(rep))
    out = subprocess.run([sys.executable, "scripts/ncu_summary.py", rep, "--stalls"], capture_output=True, text=True).stdout
    src = subprocess.run([sys.executable, "scripts/ncu_source.py", rep], capture_output=True, text=True).stdout
    with open(f"profiles/{tag}_{name}_ncu_full.txt", "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on ({rep}); read with scripts/ncu_summary.py / ncu_source.py\n")
        f.write(out); f.write("\n# per-instruction view of the first captured launch\n"); f.write(src)
print(os.listdir("profiles"))
