#!/usr/bin/env python
"""Device time of the non-sweep kernels of the step at a given size (CUDA events on the solver stream,
10 calls each after a warm-up) against their algorithmic bytes and the measured HBM peak."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--w", type=int, default=256); ap.add_argument("--h", type=int, default=256)
    ap.add_argument("--d", type=int, default=256); ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    from fluid_simulation_b200 import Simulation, masks
    from fluid_simulation_b200.simulation import OP_COPY
    sim = Simulation(a.w, a.h, a.d, 1, acc=2)
    sim.set("obs", masks.airfoil(a.w, a.h, a.d))
    sim.run_iterations(3)                       # a developed flow: realistic back-traces for advect
    pk = os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")
    peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6650.0
    cells = a.w * a.h * a.d
    ops = [("divergence", 20, lambda: sim.divergence()), ("gradient", 28, lambda: sim.gradient()),
           ("advect v_x", 16, lambda: sim.advect(1, "v_x", "v_x_prev")), ("advect v_z", 16, lambda: sim.advect(3, "v_z", "v_z_prev")),
           ("advect v_x+v_y+v_z (one launch)", 32, lambda: sim.advect_velocity()),
           ("advect dens", 20, lambda: sim.advect(0, "dens", "buffer")), ("setBounds", 8, lambda: sim.setBounds(1, "v_x")),
           ("copy", 8, lambda: sim._call(OP_COPY, 0, "buffer", "dens")), ("stats", 16, lambda: sim.stats())]
    out = {}
    for name, bpc, fn in ops:
        fn(); sim.sync()
        sim.timer_start()
        for _ in range(a.reps): fn()
        us = sim.timer_stop() * 1e3 / a.reps
        out[name] = {"us": round(us, 1), "GBps": round(bpc * cells / us / 1e3, 1), "frac": round(bpc * cells / us / 1e3 / peak, 3)}
        print(f"{name:32s} {us:8.1f} us  {bpc:2d} B/cell  {bpc * cells / us / 1e3:7.1f} GB/s  {bpc * cells / us / 1e3 / peak:5.2f} of measured peak")
    print(json.dumps({"kernels": out, "cfg": vars(a)}))
    sim.close()


if __name__ == "__main__":
    main()
