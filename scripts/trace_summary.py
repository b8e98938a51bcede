#!/usr/bin/env python
"""Summarise the per-CTA timelines FLUID_SW_TRACE writes (sweep_stream.cuh): for each file the spread of the CTAs'
start and end times and, in z-slabs, what the CTAs that begin with a face segment cost.
    python scripts/trace_summary.py gpurun_out/r2s/trace_*.txt"""
import sys

import numpy as np


def q(v):
    return " ".join(f"{x:6.1f}" for x in np.percentile(v, [0, 10, 50, 90, 100]) / 1e3)


for path in sys.argv[1:]:
    lines = open(path).read().splitlines()
    rows = [[int(x) for x in ln.split()] for ln in lines[1:]]
    a = np.array([r for r in rows if r[2] < 10 ** 9 and r[3] < 10 ** 9], dtype=np.int64)
    start, first_done, end, mode = a[:, 1], a[:, 2], a[:, 3], a[:, 4]
    face = (mode >> 1) > 0
    print(f"{path.split('/')[-1]}: {lines[0][2:].split(':')[0]}  ({len(a)} of {len(rows)} CTAs had work)")
    print(f"    us, percentiles 0 10 50 90 100      start {q(start)}    end {q(end)}")
    if face.any():
        late = ~face & (start > 2000)
        prompt = ~face & ~late
        print(f"    {face.sum():3d} CTAs begin with a face segment: it is done at {q(first_done[face])}; they end at {q(end[face])}")
        if late.any():
            print(f"    {late.sum():3d} CTAs start behind an edge CTA:  start {q(start[late])}; end {q(end[late])}")
        print(f"    {prompt.sum():3d} other CTAs:                      end {q(end[prompt])}")
