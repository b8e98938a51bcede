#!/usr/bin/env python
"""Wall time of the viewers' arithmetic on the device against the CPU checker:
streamlines with the viewer's defaults and one colour-mapped z-slice, on a developed flow."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--w", type=int, default=256); ap.add_argument("--h", type=int, default=256)
    ap.add_argument("--d", type=int, default=256); ap.add_argument("--density", type=int, default=30)
    ap.add_argument("--proximity", type=float, default=2.0)
    a = ap.parse_args()
    from fluid_simulation_b200 import Simulation, masks
    from oracle import postproc as opp
    sim = Simulation(a.w, a.h, a.d, 1, acc=10)
    sim.set("obs", masks.sphere(a.w, a.h, a.d))
    sim.run_iterations(4)
    st = {k: sim.get(k) for k in ("v_x", "v_y", "v_z", "obs", "dens")}
    out = {"cfg": vars(a)}
    for name, gpu, cpu in (
            ("streamlines", lambda: sim.streamlines(density=a.density, proximity=a.proximity),
             lambda: opp.streamlines(st["v_x"], st["v_y"], st["v_z"], st["obs"], density=a.density, proximity=a.proximity)),
            ("slice", lambda: sim.render_slice("dens", a.d // 2), lambda: opp.render_slice(st["dens"], st["obs"], a.d // 2, 0.0, 0.01))):
        gpu(); sim.sync()
        t0 = time.perf_counter(); g = gpu(); t1 = time.perf_counter(); c = cpu(); t2 = time.perf_counter()
        if name == "streamlines":
            same = [len(x) for x in g[0]] == [len(x) for x in c[0]] and all(np.array_equal(x, y) for x, y in zip(g[0], c[0]))
            out[name] = {"gpu_ms": round((t1 - t0) * 1e3, 2), "cpu_checker_ms": round((t2 - t1) * 1e3, 2), "lines": len(g[0]),
                         "points": int(sum(len(x) for x in g[0])), "bit_identical": bool(same)}
        else:
            out[name] = {"gpu_ms": round((t1 - t0) * 1e3, 2), "cpu_checker_ms": round((t2 - t1) * 1e3, 2), "bit_identical": bool(np.array_equal(g, c))}
    print(json.dumps(out))
    sim.close()


if __name__ == "__main__":
    main()
