#!/usr/bin/env python
"""bench.py -- cell-updates/s per full step (GLUPS) of the wind-tunnel hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one iteration of the reference's run() loop body (simulation.cpp:65-71): inlet
density, buffer = dens, Simulation::step() with all 6*acc solver sweeps (faithful mode: the dead
density diffusion is executed too), no frame I/O.  N = 1 workload = BASELINE.json configs[2]:
256^3 tunnel, airfoil obstacle, acc = 80 (the headline single-GPU roofline case).  N > 1 keeps
256 z-planes per GPU (weak scaling: 256 x 256 x 256N, z-slabs, halo exchange over NVLink).

Prints ONE JSON line (rank 0).  `value` is timed with CUDA events on the solver's stream with the
state resident in HBM; `e2e` is the same metric through the reference-facing API with HOST
buffers: initial obstacle upload H2D, every step followed by the reference's frame dump
(dens, obs, v_x, v_y, v_z D2H into pinned memory on a side stream), all inside the timed region.
"""
from __future__ import annotations

import argparse
import contextlib
import io
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cell_updates_per_s_per_full_step"
UNIT = "GLUPS"
W0, H0, D0, ACC0 = 256, 256, 256, 80          # BASELINE.json configs[2]


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons while the timed region runs: NVML polled every few ms (the region is
    ~0.1 s long, too short for nvidia-smi's process start-up), nvidia-smi as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu=0, uuid=None):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.stop_flag, self.source = gpu, [], False, "nvidia-smi"
        self.nv = self.dev = None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                self.dev = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode()) if uuid else None
            except Exception:
                self.dev = None
            if self.dev is None:
                self.dev = pynvml.nvmlDeviceGetHandleByIndex(gpu)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM))
            self.nv, self.source = pynvml, "nvml"
        except Exception:
            self.nv = None

    def sample_nvml(self):
        nv = self.nv
        mhz = nv.nvmlDeviceGetClockInfo(self.dev, nv.NVML_CLOCK_SM)
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
            else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
        bits = [nv.nvmlClocksThrottleReasonHwSlowdown, nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                nv.nvmlClocksThrottleReasonSwThermalSlowdown, nv.nvmlClocksThrottleReasonSwPowerCap]
        return [str(mhz), str(self.max_mhz)] + ["Active" if r & b else "Not Active" for b in bits]

    def run(self):
        while not self.stop_flag:
            try:
                if self.nv is not None:
                    self.samples.append(self.sample_nvml())
                    time.sleep(0.004)
                    continue
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.samples.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        sm = [float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        reasons = sorted({n for s in self.samples for n, v in zip(self.NAMES, s[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": self.source}


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one single-field k_sweep_stream launch at 256^3, from
    the newest committed `ncu --set full` summary under profiles/ (bytes; None if there is none)."""
    import glob
    import re
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full.txt")), reverse=True):
        txt = open(path).read()
        rd = re.search(r"dram__bytes_read\.sum = ([0-9.]+) Mbyte", txt)
        wr = re.search(r"dram__bytes_write\.sum = ([0-9.]+) Mbyte", txt)
        if rd and wr and "k_sweep_stream" in txt:
            return (float(rd.group(1)) + float(wr.group(1))) * 1e6, os.path.relpath(path, ROOT)
    return None, None


def workload(n_gpus, args):
    w, h = args.w or W0, args.h or H0
    d = (args.d or D0) * (1 if getattr(args, "strong", False) else n_gpus)     # --strong: --d is the whole tunnel
    return w, h, d, (args.acc or ACC0)


def bytes_per_cell_step(acc, elided=False):
    # B_faithful(K) = 196 + 72 K, B_elided(K) = 196 + 60 K
    return 196 + (60 if elided else 72) * acc


# ------------------------------------------------------------------------------------------------
def cpu_reference_sample(w, h, d, acc, mask, threads=None, max_seconds=40.0):
    """Time the reference's own CPU code (oracle/_ref, unmodified object code, all host threads) on a
    bounded sample of the workload: full grid, run() iterations at acc = 1 and acc = 3; the per-step
    time is linear in acc (6*acc sweeps dominate), so t(acc) = t0 + acc*ts is extrapolated to the
    config's acc.  Falls back to the plain-C port when the reference build did not travel."""
    from oracle import RefSim, OracleSim, have_ref
    kind = "reference" if have_ref() else "port"
    if kind == "reference":
        threads = threads or RefSim.max_threads()
        mk = lambda a: RefSim(w, h, d, acc=a, threads=threads)
    else:
        threads = 1
        mk = lambda a: OracleSim(w, h, d, acc=a)
    times = {}
    for a in (1, 3):
        sim = mk(a)
        sim.set("obs", mask)
        t = time.perf_counter()
        sim.run_iterations(1)
        times[a] = time.perf_counter() - t
        del sim
    ts = max((times[3] - times[1]) / 2.0, 1e-9)
    t0 = max(times[1] - ts, 0.0)
    t_full = t0 + acc * ts
    glups = w * h * d / t_full / 1e9
    sample = (f"{w}x{h}x{d} full grid, 1 run() iteration at acc=1 ({times[1]:.2f} s) and acc=3 "
              f"({times[3]:.2f} s) on {threads} thread(s); per-step time extrapolated linearly in acc "
              f"to acc={acc}: {t_full:.1f} s/step")
    return {"value": glups, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample,
            "s_per_step_extrapolated": t_full}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from fluid_simulation_b200 import masks
    w, h, d, acc = workload(args.gpus, args)
    mask = masks.airfoil(w, h, d)
    vals = []
    budget = time.time() + 200.0
    n_run = 0
    for i in range(args.warmup + args.steps):
        if i >= 1 + min(args.warmup, 1) and time.time() > budget:
            break
        r = cpu_reference_sample(w, h, d, acc, mask)
        if i >= min(args.warmup, 1):
            vals.append(r)
        n_run += 1
    best = max(vals, key=lambda r: r["value"])
    mean_s = float(np.mean([r["s_per_step_extrapolated"] for r in vals]))
    value = w * h * d / mean_s / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": mean_s * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{w}x{h}x{d} tunnel, airfoil obstacle, acc={acc} (BASELINE.json configs[2])",
                   "parallelism": f"OpenMP x{best['cores']} on host cores",
                   "samples_executed": len(vals)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": best["cores"], "kind": best["kind"],
                         "sample": best["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    n = args.gpus
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from fluid_simulation_b200 import Simulation, masks, SOLVER_RB, SOLVER_JACOBI, ELIDE_DEAD, _lib
    from fluid_simulation_b200 import slab as slabmod

    w, h, d, acc = workload(world, args)
    flags = {"rb": SOLVER_RB, "jacobi": SOLVER_JACOBI}[args.solver] | (ELIDE_DEAD if args.elide else 0)
    mask = masks.airfoil(w, h, d)
    L = _lib.load()
    L.fluid_set_device(local)
    sim = Simulation(w, h, d, args.steps, acc=acc, flags=flags, rank=rank, nranks=world)
    if world > 1:
        slabmod.attach_nccl(sim, dist)
    sim.set("obs", mask)

    def barrier():
        sim.sync()
        if dist is not None:
            dist.barrier()

    def reduce_max(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident timing ------------------------------------------------------------
    sim.run_iterations(args.warmup)
    # everything slow (NVML start-up is tens of ms) happens BEFORE the barrier: a rank that enters the timed
    # region late keeps its z-neighbours waiting, and the reported time is the maximum over the ranks
    vis = [v.strip() for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v.strip()]
    entry = vis[local] if local < len(vis) else str(local)          # NVML numbers physical devices
    sampler = (ClockSampler(int(entry)) if entry.isdigit() else ClockSampler(local, entry[4:] if entry.startswith("GPU-") else None)) \
        if rank == 0 else None
    sim.profile(True); sim.profile_read()
    barrier()
    if sampler:
        sampler.start()
    l0 = sim.launch_count()
    sim.timer_start()
    sim.run_iterations(args.steps)
    ms = sim.timer_stop()
    barrier()
    launches = sim.launch_count() - l0
    solve_ms, field_sweeps, _ = sim.profile_read()
    sim.profile(False)
    ms = reduce_max(ms)
    clocks = sampler.summary() if sampler else None
    cells = w * h * d
    value = cells * args.steps / (ms * 1e-3) / 1e9
    stats = sim.stats()

    # ---- roofline of the dominant kernel family (solver sweeps) ------------------------------
    peak, peak_src = peaks()
    local_cells = w * h * sim.d_local
    sweep_bytes = 12.0 * local_cells                      # x in, rhs in, x out per cell per sweep
    achieved = sweep_bytes * field_sweeps / (solve_ms * 1e-3) / 1e9 if solve_ms > 0 else 0.0
    step_bytes = bytes_per_cell_step(acc, args.elide) * local_cells
    traffic, traffic_src = ncu_traffic()
    roofline = {
        "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic if (w, h, sim.d_local) == (W0, H0, D0) else None,
        "traffic_source": f"{traffic_src}: dram read+write bytes of one single-field sweep launch "
                          f"(algorithmic {sweep_bytes / 1e6:.1f} MB)" if traffic_src else None,
        "kernel": "linearSolver sweeps (12 B/cell/sweep algorithmic; CUDA events around every solve)",
        "peak_source": peak_src, "us_per_field_sweep": solve_ms * 1e3 / max(field_sweeps, 1),
        "solver_share_of_step": solve_ms / (ms if ms > 0 else 1.0),
        "whole_step": {"algorithmic_bytes": step_bytes,
                       "achieved": step_bytes * args.steps / (ms * 1e-3) / 1e9,
                       "frac": step_bytes * args.steps / (ms * 1e-3) / 1e9 / peak,
                       "frac_of_nominal_8000": step_bytes * args.steps / (ms * 1e-3) / 1e9 / 8000.0},
    }

    # ---- end to end through the reference-facing API with HOST buffers -----------------------
    e2e = run_e2e(sim, args, w, h, d, mask, barrier, reduce_max, rank, world)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            cpu = cpu_reference_sample(w, h, d, acc, mask)
            cpu.pop("s_per_step_extrapolated", None)
        except Exception as ex:  # the checker is optional for the measurement itself
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(ex)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{w}x{h}x{d} tunnel, airfoil obstacle, acc={acc} "
                                   + ("(BASELINE.json configs[2]" + (f" x{world} z-slabs" if world > 1 else "") + ")"
                                      if (w, h, acc, args.strong) == (W0, H0, ACC0, False) else "(custom size)"),
                       "solver": args.solver, "density_diffusion": "elided" if args.elide else "executed",
                       "parallelism": f"z-slab x{world}" if world > 1 else "1 GPU",
                       "l2": "state (11 fields, %.0f MB) exceeds the 126 MB L2" % (11 * 4 * local_cells / 1e6)},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
            "cpu_baseline": cpu,
            "stats": {k: float(v) for k, v in stats.items()},
        }
        print(json.dumps(line), flush=True)
    sim.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def run_e2e(sim, args, w, h, d, mask, barrier, reduce_max, rank, world):
    """run() as a host program would drive it: the inlet density is injected from HOST arrays
    through the public mutator (simulation.cpp:65-67 calls addDensity per face cell), then
    buffer = dens, step(), and the frame write of step() (simulation.cpp:140-148) as an async D2H
    of dens, v_x, v_y, v_z into pinned host memory (obs is constant and already host-resident)."""
    from fluid_simulation_b200.simulation import OP_STEP, OP_COPY
    zs = np.arange(sim.z_lo, sim.z_lo + sim.d_local, dtype=np.int32)
    ys = np.arange(1, h + 1, dtype=np.int32)
    zz, yy = np.meshgrid(zs, ys, indexing="ij")
    xyz = np.stack([np.ones_like(yy), yy, zz], axis=-1).reshape(-1, 3).astype(np.int32)
    xyz_p = sim.pinned_array(xyz.shape, np.int32); xyz_p[...] = xyz
    amt_p = sim.pinned_array((len(xyz),), np.float32); amt_p[...] = 0.001
    steps = max(2, min(args.steps, 10))
    # one-time setup outside the timed region, like the reference opening its files: the pinned frame
    # buffers (cudaHostAlloc of 4 x 69 MB takes longer than a whole step) and the device staging
    sim.dump_frame_async((True, False, True, True, True)); sim.frame_wait()
    barrier()
    t0 = time.perf_counter()
    sim.set("obs", mask)                                  # obstacle mask H2D (once, amortised)
    for _ in range(steps):
        sim.add_density_cells(xyz_p, amt_p)               # H2D: this step's inlet source
        sim._call(OP_COPY, 0, "buffer", "dens")           # buffer = dens   (simulation.cpp:70)
        sim._call(OP_STEP)                                # step()          (simulation.cpp:71)
        sim.dump_frame_async((True, False, True, True, True))   # frame D2H on the side stream
    sim.frame_wait()
    barrier()
    dt = reduce_max(time.perf_counter() - t0)
    plane = (w + 2) * (h + 2) * 4
    d2h = 4 * plane * (sim.d_local + (2 if world == 1 else 1))
    h2d = xyz_p.nbytes + amt_p.nbytes + mask.nbytes // world // steps
    return {"value": w * h * d * steps / dt / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "steps": steps, "ms_per_step": dt * 1e3 / steps,
            "path": "fluid_add_density(H2D) + buffer=dens + step() + fluid_dump_frame_async(D2H, pinned)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--solver", default="rb", choices=["rb", "jacobi"])
    ap.add_argument("--elide", action="store_true", help="skip the dead density diffusion (F2)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--w", type=int); ap.add_argument("--h", type=int); ap.add_argument("--d", type=int)
    ap.add_argument("--acc", type=int)
    ap.add_argument("--size", help="WxHxD, same as --w/--h/--d (whose short names torchrun's own parser trips over)")
    ap.add_argument("--strong", action="store_true", help="--d is the depth of the whole tunnel, split over the GPUs "
                    "(BASELINE.json configs[3], configs[4]); default: --d planes per GPU (weak scaling)")
    args = ap.parse_args()
    if args.size:
        args.w, args.h, args.d = (int(v) for v in args.size.lower().split("x"))
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3                                   # timing rule: W >= 3
    # stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner) are sent
    # to stderr for the duration of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    out = io.StringIO()
    try:
        with contextlib.redirect_stdout(out):
            rc = run_reference_arm(args) if args.impl == "reference" else run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    lines = [l for l in out.getvalue().splitlines() if l.startswith("{")]
    if lines:
        print(lines[-1], flush=True)
    return rc


if __name__ == "__main__":
    sys.exit(main())
