#!/usr/bin/env python
"""Where does the end-to-end step time go?  D2H / H2D bandwidth of pinned buffers and the e2e loop of
bench.py with its pieces switched off one at a time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from fluid_simulation_b200 import Simulation, masks
from fluid_simulation_b200.simulation import OP_STEP, OP_COPY

n = 69 * 1024 * 1024
d = torch.empty(n, dtype=torch.uint8, device="cuda"); hp = torch.empty(n, dtype=torch.uint8).pin_memory()
for name, fn in (("D2H pinned", lambda: hp.copy_(d, non_blocking=True)), ("H2D pinned", lambda: d.copy_(hp, non_blocking=True))):
    fn(); torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(5): fn()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 5
    print(f"{name}: {n / dt / 1e9:.1f} GB/s")

w = h = dd = 256
sim = Simulation(w, h, dd, 1, acc=80)
sim.set("obs", masks.airfoil(w, h, dd))
sim.run_iterations(2); sim.sync()
def loop(steps, dump, which=(True, False, True, True, True)):
    sim.sync(); t = time.perf_counter()
    for _ in range(steps):
        sim._call(OP_COPY, 0, "buffer", "dens"); sim._call(OP_STEP)
        if dump: sim.dump_frame_async(which)
    if dump: sim.frame_wait()
    sim.sync(); return (time.perf_counter() - t) / steps * 1e3
print("no dump      %.1f ms/step" % loop(4, False))
print("dump 4 fields %.1f ms/step" % loop(4, True))
print("dump 1 field  %.1f ms/step" % loop(4, True, (True, False, False, False, False)))
t = time.perf_counter(); sim.dump_frame_async((True, False, True, True, True)); sim.frame_wait(); print("one dump alone %.1f ms" % ((time.perf_counter() - t) * 1e3))
