#!/usr/bin/env python
"""Timing of the CUDA STL voxeliser against the reference's own loader on the host cores.
    python scripts/voxel_bench.py [--w 256 --h 256 --d 256 --lat 48 --lon 96]"""
import argparse, json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--w", type=int, default=256); ap.add_argument("--h", type=int, default=256)
    ap.add_argument("--d", type=int, default=256); ap.add_argument("--lat", type=int, default=48)
    ap.add_argument("--lon", type=int, default=96); ap.add_argument("--no-ref", action="store_true")
    a = ap.parse_args()
    from fluid_simulation_b200 import Simulation, stl
    tris = stl.sphere_mesh(1.0, a.lat, a.lon)
    out = {"triangles": len(tris), "grid": [a.w, a.h, a.d]}
    sim = Simulation(a.w, a.h, a.d, 1)
    sim.voxelize(tris, 0.8)                                   # warm-up (module load, allocations)
    for mode, name in ((0, "hashed"), (1, "fixed"), (2, "cells")):
        sim.set("obs", np.zeros(sim.shape, np.float32))
        sim.sync(); t = time.perf_counter()
        added = sim.voxelize(tris, 0.8, (10.0, 20.0, 30.0), (0.0, 0.0, 0.0), mode)
        dt = time.perf_counter() - t
        out[name] = {"ms": dt * 1e3, "added": added, "cells": int((sim.get("obs") > 0).sum())}
    if not a.no_ref:
        import oracle
        if oracle.have_ref():
            p = os.path.join(tempfile.mkdtemp(), "s.stl"); stl.write_stl_binary(p, tris)
            t = time.perf_counter()
            ref = oracle.ref_voxelize(p, a.w, a.h, a.d, 0.8, (10.0, 20.0, 30.0), (0.0, 0.0, 0.0))
            out["reference_cpu"] = {"ms": (time.perf_counter() - t) * 1e3, "cells": int((ref > 0).sum()),
                                    "threads": oracle.RefSim.max_threads()}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
