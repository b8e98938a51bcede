"""Worker of the multi-GPU parity tests: run under `python -m torch.distributed.run --nproc-per-node N`.
Each rank drives one z-slab; rank 0 assembles the owned planes of every rank and compares the global
fields bit-for-bit with the single-domain CPU oracle.  Prints `SLAB-PARITY OK <case>` per case."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from fluid_simulation_b200 import Simulation, masks, _lib, multi_gpu, SOLVER_RB, SOLVER_JACOBI, FIELD_NAMES
    from fluid_simulation_b200 import slab as slabmod
    _lib.load().fluid_set_device(local)
    from oracle import OracleSim, ORDER_RB, ORDER_JACOBI

    def gather_global(sim):
        return multi_gpu.gather_global_state(sim, dist, FIELD_NAMES)

    failures = 0
    cases = [("rb", SOLVER_RB, ORDER_RB), ("jacobi", SOLVER_JACOBI, ORDER_JACOBI)]
    # thin slabs (whole-column sweeps + hand-over kernel), thick ragged slabs (edge CTAs hand the halos over
    # inside the sweep), odd sweep counts (the solver's buffers end up permuted, on every rank alike), and the
    # four-warp-column kernel (256 < W <= 512) and the x-tiled two-column kernel (W > 512: 516 -> 3 tiles, 772 -> 4)
    # (every slab keeps at least FL_HALO = 4 planes)
    shapes = [(24, 16, max(12, 4 * world + 1), 4), (128, 30, max(17, 5 * world), 4), (130, 9, max(9, 4 * world), 4), (256, 40, 4 * world + 3, 4), (132, 20, 35 * world + 1, 4),
              (132, 18, 33 * world, 3), (260, 12, 33 * world + 2, 3), (516, 6, 32 * world, 4), (772, 10, 34 * world, 3)]
    for name, flag, order in cases:
        for (w, h, d, acc) in shapes:
            st = masks.random_state(w, h, d, seed=11)
            gpu = Simulation(w, h, d, 1, acc=acc, flags=flag, rank=rank, nranks=world)
            multi_gpu.attach(gpu, dist)
            gpu.load_state(st)
            cpu = OracleSim(w, h, d, acc=acc, order=order) if rank == 0 else None
            if cpu: cpu.load_state(st)
            steps = [("linearSolver", lambda s: s.linearSolver(3, "v_z", "v_z_prev", 0.37, 3.22),
                      lambda o: o.linear_solver(3, "v_z", "v_z_prev", 0.37, 3.22)),
                     ("diffuse", lambda s: s.diffuse(0, "dens", "buffer"), lambda o: o.diffuse(0, "dens", "buffer")),
                     ("project", lambda s: s.project(), lambda o: o.project()),
                     ("advect1", lambda s: s.advect(1, "v_x", "v_x_prev"), lambda o: o.advect(1, "v_x", "v_x_prev")),
                     ("advect0", lambda s: s.advect(0, "dens", "buffer"), lambda o: o.advect(0, "dens", "buffer")),
                     ("inflow", lambda s: (s.inflow_density(), s.inflow_velocity()),
                      lambda o: (o.inflow_density(), o.inflow_velocity())),
                     ("step", lambda s: s.step(), lambda o: o.step()),
                     ("run3", lambda s: s.run_iterations(3), lambda o: o.run_iterations(3))]
            for tag, fg, fc in steps:
                fg(gpu)
                got = gather_global(gpu)
                if rank == 0:
                    fc(cpu)
                    want = cpu.state()
                    bad = [n for n in FIELD_NAMES
                           if not np.array_equal(got[n].view(np.uint32), want[n].view(np.uint32))]
                    case = f"{name} {w}x{h}x{d} x{world} {tag}"
                    if bad:
                        failures += 1
                        for n in bad:
                            dd = np.abs(got[n].astype(np.float64) - want[n])
                            print(f"SLAB-PARITY FAIL {case}: {n} {np.count_nonzero(got[n] != want[n])} cells, "
                                  f"max {dd.max():.3e} at {np.unravel_index(np.argmax(dd), dd.shape)}", flush=True)
                    else:
                        print(f"SLAB-PARITY OK {case}", flush=True)
            gpu.close()
            dist.barrier()
    t = torch.tensor([failures], device="cuda")
    dist.broadcast(t, 0)
    dist.destroy_process_group()
    sys.exit(1 if int(t.item()) else 0)


if __name__ == "__main__":
    main()
