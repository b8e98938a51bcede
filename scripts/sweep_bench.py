#!/usr/bin/env python
"""Micro-benchmark of the linearSolver sweep kernels alone (device-resident, CUDA events around each
solve through fluid_profile_*): us per field-sweep and fraction of the HBM roofline at 12 B/cell.
    python scripts/sweep_bench.py [--w 256 --h 256 --d 256 --acc 40 --reps 3 --solver rb]
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--w", type=int, default=256); ap.add_argument("--h", type=int, default=256)
    ap.add_argument("--d", type=int, default=256); ap.add_argument("--acc", type=int, default=40)
    ap.add_argument("--reps", type=int, default=3); ap.add_argument("--solver", default="rb")
    ap.add_argument("--flags", type=lambda s: int(s, 0), default=0)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    from fluid_simulation_b200 import Simulation, masks, SOLVER_RB, SOLVER_JACOBI
    flag = {"rb": SOLVER_RB, "jacobi": SOLVER_JACOBI}[a.solver] | a.flags
    sim = Simulation(a.w, a.h, a.d, 1, acc=a.acc, flags=flag)
    sim.set("obs", masks.airfoil(a.w, a.h, a.d))
    rng = np.random.default_rng(0)
    sim.set("divergence", rng.standard_normal(sim.shape).astype(np.float32))
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0
    out = {}
    # single field (pressure solve) and the 3-field batch (velocity diffusion, through step())
    sim.linearSolver(0, "pressure", "divergence", 1.0, 6.0)          # warm-up
    sim.profile(True); sim.profile_read()
    for _ in range(a.reps):
        sim.linearSolver(0, "pressure", "divergence", 1.0, 6.0)
    ms, sweeps, calls = sim.profile_read()
    us1 = ms * 1e3 / sweeps
    cells = a.w * a.h * a.d
    out["single"] = {"us_per_sweep": us1, "GBps": 12.0 * cells / us1 / 1e3, "frac": 12.0 * cells / us1 / 1e3 / peak}
    sim.profile(False)
    sim.run_iterations(1)
    sim.profile(True); sim.profile_read()
    sim.run_iterations(a.reps)
    ms, sweeps, calls = sim.profile_read()
    us = ms * 1e3 / sweeps
    out["step_mix"] = {"us_per_field_sweep": us, "GBps": 12.0 * cells / us / 1e3, "frac": 12.0 * cells / us / 1e3 / peak}
    out["cfg"] = {"w": a.w, "h": a.h, "d": a.d, "acc": a.acc, "solver": a.solver, "tag": a.tag,
                  "env": {k: v for k, v in os.environ.items() if k.startswith("FLUID_")}}
    print(json.dumps(out))
    sim.close()


if __name__ == "__main__":
    main()
