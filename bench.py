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
def host_threads():
    """Host cores this process may use.  NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_reference_sample(w, h, d, acc, mask, threads=None, iterations=1, sim=None):
    """Time the reference's own CPU code (oracle/_ref: unmodified object code, all host threads) on the
    workload itself: the full grid at the config's acc, `iterations` real run() iterations (inlet, buffer,
    step() with all 6*acc sweeps; no frame I/O) after one untimed acc=1 iteration that takes the first-touch
    page faults and starts the OpenMP pool.  Nothing is extrapolated.  Falls back to the plain-C port (one
    thread) when the reference build did not travel.  Returns the sample and the warmed simulation."""
    from oracle import RefSim, OracleSim, have_ref
    kind = "reference" if have_ref() else "port"
    if sim is None:
        if kind == "reference":
            threads = threads or host_threads()
            sim = RefSim(w, h, d, acc=1, threads=threads)
        else:
            threads = 1
            sim = OracleSim(w, h, d, acc=1)
        sim.set("obs", mask)
        sim.run_iterations(1)                              # untimed warm-up (acc = 1)
        sim.set_acc(acc)
    elif kind == "reference":
        threads = threads or host_threads()
        RefSim.set_threads(threads)
    else:
        threads = 1
    t = time.perf_counter()
    sim.run_iterations(iterations)
    s_per_step = (time.perf_counter() - t) / iterations
    glups = w * h * d / s_per_step / 1e9
    sample = (f"{w}x{h}x{d} full grid, {iterations} real run() iteration(s) at acc={acc} on {threads} thread(s) after one "
              f"untimed acc=1 iteration: {s_per_step:.2f} s/step (measured, not extrapolated)")
    return {"value": glups, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample,
            "s_per_step": s_per_step}, sim


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation on the box's host cores.  Every timed "step" is one
    real run() iteration at the config's acc.  N = 1: BASELINE.json configs[2] itself (256^3).  N > 1: the CPU is
    not run above 256^3 (hours per step), so the arm times ONE GPU's 256-plane slab of the weak-scaled
    tunnel and says so; its cells/s are size-independent to first order."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from fluid_simulation_b200 import masks
    w, h, d_total, acc = workload(args.gpus, args)
    d = d_total if (args.gpus == 1 or args.strong) and d_total <= 256 else (args.d or D0)
    mask = masks.airfoil(w, h, d)
    budget = time.time() + float(os.environ.get("FLUID_REF_BUDGET_S", "150"))
    per_step, sim, first = [], None, None
    for i in range(args.warmup + args.steps):
        r, sim = cpu_reference_sample(w, h, d, acc, mask, sim=sim)
        first = first or r
        if i >= min(args.warmup, 1) or args.warmup + args.steps == 1:
            per_step.append(r["s_per_step"])
        if per_step and time.time() + r["s_per_step"] > budget:
            break                                          # bounded: a real step is ~25 s on 16 threads
    mean_s = float(np.mean(per_step))
    value = w * h * d / mean_s / 1e9
    sub = "" if d == d_total else f"; CPU timed on one GPU's {w}x{h}x{d} slab of the {w}x{h}x{d_total} tunnel (not run on CPU above 256^3)"
    sample = (f"{w}x{h}x{d} full grid, {len(per_step)} real run() iteration(s) at acc={acc} on {first['cores']} thread(s), "
              f"mean {mean_s:.2f} s/step (min {min(per_step):.2f}, max {max(per_step):.2f}); one untimed acc=1 iteration and "
              f"{min(args.warmup, 1)} untimed real iteration(s) first; measured, not extrapolated")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(per_step), "steps_requested": args.steps, "warmup": args.warmup, "ms_per_step": mean_s * 1e3,
        "higher_is_better": True,
        "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{w}x{h}x{d_total} tunnel, airfoil obstacle, acc={acc} (BASELINE.json configs[2]"
                               + (f" x{args.gpus} z-slabs" if args.gpus > 1 else "") + ")" + sub,
                   "parallelism": f"OpenMP x{first['cores']} on host cores",
                   "samples_executed": len(per_step)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": first["cores"], "kind": first["kind"], "sample": sample},
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
    bound = None
    if world > 1:                       # node-local pinned frame buffers: each rank next to its own GPU
        from fluid_simulation_b200 import multi_gpu
        bound = multi_gpu.bind_host_to_gpu(local)

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

    # ---- multi-GPU parity where the driver sees it (outside every timed region) ---------------
    slab_parity = run_slab_parity(rank, world, dist) if world > 1 else None

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

    # ---- the other configurations north_star names, measured in the same (driver-run) process -----
    extra = None
    if not args.no_extra and (w, h, acc, args.strong) == (W0, H0, ACC0, False):
        sim.close()
        extra = run_extra(rank, world, dist, local)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            cpu, _cpu_sim = cpu_reference_sample(w, h, d, acc, mask)
            del _cpu_sim
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
                       "l2": "state (11 fields, %.0f MB) exceeds the 126 MB L2" % (11 * 4 * local_cells / 1e6),
                       "host_affinity": (f"each rank pinned to the {len(bound)} CPUs NVML lists for its GPU" if bound
                                         else "unchanged")},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
            "cpu_baseline": cpu, "slab_parity": slab_parity, "extra": extra,
            "stats": {k: float(v) for k, v in stats.items()},
        }
        print(json.dumps(line), flush=True)
    if extra is None:
        sim.close()
    if dist is not None:
        dist.destroy_process_group()
    if slab_parity is not None and not slab_parity["bit_identical"]:
        return 3                                          # a fast wrong answer is not a result
    return 0


def run_extra(rank, world, dist, local):
    """BASELINE.json's scaling configurations, device-timed like `value` (CUDA events, max over ranks), a few
    steps each so that the whole bench still ends within minutes:
      configs[3]  512^3 tunnel, acc 80, split over the N ranks (strong scaling; N = 1 is the base of the curve)
      configs[4]  1024 x 512 x 512, acc 200, 8 ranks, with the frame dump every 10 steps (pinned async D2H)
      target      1024^3, acc 80, 8 ranks, beside 1024 x 1024 x 128 on one GPU (the base of its efficiency)."""
    from fluid_simulation_b200 import Simulation, masks, multi_gpu, SOLVER_RB
    out = {}

    def timed(tag, w, h, d, acc, steps, warm, dump_every=0):
        sim = Simulation(w, h, d, steps, acc=acc, flags=SOLVER_RB, rank=rank, nranks=world)
        if world > 1:
            multi_gpu.attach(sim, dist)
        sim.set("obs", masks.airfoil_planes(w, h, d, sim.z_lo - 4, sim.d_local + 8))
        sim.run_iterations(warm)
        if dump_every:
            sim.dump_frame_async(); sim.frame_wait()       # pinned buffers exist before the clock starts
        sim.sync()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        sim.timer_start()
        for i in range(steps):
            sim.run_iterations(1)
            if dump_every and (i + 1) % dump_every == 0:
                sim.dump_frame_async()
        ms = sim.timer_stop()
        if dump_every:
            sim.frame_wait()
        wall = (time.perf_counter() - t0) * 1e3
        if dist is not None:
            import torch
            t = torch.tensor([ms, wall], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1])
        sim.close()
        r = {"grid": f"{w}x{h}x{d}", "acc": acc, "n_gpus": world, "steps": steps, "ms_per_step": ms / steps,
             "glups": w * h * d * steps / (ms * 1e-3) / 1e9}
        if dump_every:
            r.update({"dump_every": dump_every, "ms_per_step_with_dumps_wall": wall / steps,
                      "frame_bytes": 5 * 4 * (w + 2) * (h + 2) * (d + 2)})
        out[tag] = r

    timed("configs3_512cubed_strong", 512, 512, 512, 80, 3, 1)
    if world == 1:
        timed("base_1024x1024x128_one_gpu", 1024, 1024, 128, 80, 2, 1)
    if world == 8:
        timed("configs4_1024x512x512_acc200_dump10", 1024, 512, 512, 200, 10, 1, dump_every=10)
        timed("target_1024cubed", 1024, 1024, 1024, 80, 2, 1)
    return out


def run_slab_parity(rank, world, dist):
    """world > 1: before anything is timed, small z-slab cases through the same library and communicator
    set-up as the benchmark, compared bit for bit on rank 0 with the single-domain CPU oracle (ordering-matched
    red-black loop): a whole step() and three run() iterations each.  The slabs are thick enough for the fused
    sweep + halo hand-over (face CTAs), and interior ranks have two faces."""
    import torch
    from fluid_simulation_b200 import Simulation, masks, multi_gpu, SOLVER_RB, FIELD_NAMES
    cases, bad = 0, []
    for (w, h, d, acc) in [(132, 20, 35 * world + 1, 4), (256, 24, 33 * world, 3)]:
        st = masks.random_state(w, h, d, seed=11)
        gpu = Simulation(w, h, d, 1, acc=acc, flags=SOLVER_RB, rank=rank, nranks=world)
        multi_gpu.attach(gpu, dist)
        gpu.load_state(st)
        cpu = None
        if rank == 0:
            from oracle import OracleSim, ORDER_RB
            cpu = OracleSim(w, h, d, acc=acc, order=ORDER_RB)
            cpu.load_state(st)
        for tag, fg, fc in (("step", lambda s: s.step(), lambda o: o.step()),
                            ("run3", lambda s: s.run_iterations(3), lambda o: o.run_iterations(3))):
            fg(gpu)
            got = multi_gpu.gather_global_state(gpu, dist, FIELD_NAMES)
            cases += 1
            if rank == 0:
                fc(cpu)
                want = cpu.state()
                for n in FIELD_NAMES:
                    if not np.array_equal(got[n].view(np.uint32), want[n].view(np.uint32)):
                        bad.append(f"{w}x{h}x{d} {tag} {n}")
        gpu.close()
        dist.barrier()
    t = torch.tensor([len(bad)], device="cuda")
    dist.broadcast(t, 0)
    return {"ranks": world, "cases": cases, "bit_identical": int(t.item()) == 0,
            "against": "single-domain CPU oracle, ordering-matched red-black, all 11 fields", "failed": bad[:8]}


def run_e2e(sim, args, w, h, d, mask, barrier, reduce_max, rank, world):
    """run() as a host program would drive it: the inlet density is injected from HOST arrays
    through the public mutator (simulation.cpp:65-67 calls addDensity per face cell), then
    buffer = dens, step(), and the frame write of step() (simulation.cpp:140-148) as an async D2H
    of all five frame fields (dens, obs, v_x, v_y, v_z, as the reference writes them) into pinned host memory."""
    from fluid_simulation_b200.simulation import OP_STEP, OP_COPY
    zs = np.arange(sim.z_lo, sim.z_lo + sim.d_local, dtype=np.int32)
    ys = np.arange(1, h + 1, dtype=np.int32)
    zz, yy = np.meshgrid(zs, ys, indexing="ij")
    xyz = np.stack([np.ones_like(yy), yy, zz], axis=-1).reshape(-1, 3).astype(np.int32)
    xyz_p = sim.pinned_array(xyz.shape, np.int32); xyz_p[...] = xyz
    amt_p = sim.pinned_array((len(xyz),), np.float32); amt_p[...] = 0.001
    mask_p = sim.stage_field(mask)                        # pinned copy of the planes this rank holds
    steps = max(2, min(args.steps, 10))
    # one-time setup outside the timed region, like the reference opening its files: the pinned frame
    # buffers (cudaHostAlloc of 4 x 69 MB takes longer than a whole step) and the device staging
    sim.dump_frame_async(); sim.frame_wait()
    barrier()
    td = time.perf_counter()                              # one frame dump alone, every rank at once: the host link's
    sim.dump_frame_async(); sim.frame_wait()              # share of a step when the dumps do not hide behind compute
    dump_alone = reduce_max(time.perf_counter() - td)
    barrier()
    t0 = time.perf_counter()
    sim.set_staged("obs", mask_p)                         # obstacle mask H2D from pinned memory (once, amortised)
    for _ in range(steps):
        sim.add_density_cells(xyz_p, amt_p)               # H2D: this step's inlet source
        sim._call(OP_COPY, 0, "buffer", "dens")           # buffer = dens   (simulation.cpp:70)
        sim._call(OP_STEP)                                # step()          (simulation.cpp:71)
        sim.dump_frame_async()                            # all five fields, D2H on the side stream
    sim.frame_wait()
    barrier()
    dt = reduce_max(time.perf_counter() - t0)
    plane = (w + 2) * (h + 2) * 4
    d2h = 5 * plane * (d + 2)                             # whole job: every rank dumps the planes it owns
    h2d = (xyz_p.nbytes + amt_p.nbytes) * world + mask.nbytes // steps   # whole job (the ranks' halo planes of the mask not counted)
    return {"value": w * h * d * steps / dt / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "steps": steps, "ms_per_step": dt * 1e3 / steps,
            "frame_dump_alone_ms": dump_alone * 1e3, "frame_dump_alone_gbs_per_gpu": d2h / world / dump_alone / 1e9,
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
    ap.add_argument("--no-extra", action="store_true", help="skip the extra scaling configurations (512^3 strong, ...)")
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
