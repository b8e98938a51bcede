"""Attach a z-slab `Simulation` (one process per GPU) to its peers.

The solver library owns the data path: halo planes move by ncclSend/ncclRecv inside libfluid_cuda
(its own communicator) and advection gathers read the peers' slabs through CUDA IPC mappings.  This
module only does the rendezvous that needs a process group -- `torch.distributed` is plumbing here:
broadcast of the 128-byte NCCL unique id and all-gather of the 64-byte IPC handles.  It works with any
backend (`nccl` on the GPU box; the `gloo` tests exercise the same code with a fake library).
"""
from __future__ import annotations

import ctypes as C
import os

from .simulation import FIELDS, FluidError

# the fields advect() gathers from: prev_field of each of the four calls (simulation.cpp:125-127, 136)
GATHER_SOURCES = ("v_x_prev", "v_y_prev", "v_z_prev", "buffer")
# step() trades v and v_prev instead of copying (simulation.cpp:108-110), so the allocations behind v_x, v_y, v_z are
# gather sources every other step: mapped from every rank as well (without them the library keeps the copies)
GATHER_PARTNERS = ("v_x", "v_y", "v_z")
IPC_COUNT = len(FIELDS) + 5            # fields, 4 solver ping-pong partners, flag block (FLUID_IPC_COUNT)
IPC_SYNC = len(FIELDS) + 4             # FLUID_IPC_SYNC: arrival flags of the z-neighbours, barrier epochs of every rank


def nccl_library_path() -> str:
    """The libnccl the host process already uses (torch's bundled wheel), else the system one."""
    try:
        import nvidia.nccl as n
        p = os.path.join(os.path.dirname(n.__file__), "lib", "libnccl.so.2")
        if os.path.exists(p):
            return p
    except Exception:
        pass
    return "libnccl.so.2"


def bind_host_to_gpu(local_rank: int):
    """Pin this process to the CPUs NVML names as closest to its GPU (same NUMA node, same PCIe root), so that
    the pinned frame buffers it allocates afterwards are node-local (first touch) and eight ranks dumping at once
    do not all write through one socket.  Returns the CPU list, or None when NVML or the call is unavailable
    (nothing is changed then).  FLUID_NO_BIND=1 switches it off (A/B)."""
    if os.environ.get("FLUID_NO_BIND"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        index = local_rank                                   # NVML counts physical devices
        visible = os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",")
        if all(v.strip().isdigit() for v in visible) and local_rank < len(visible):
            index = int(visible[local_rank])
        handle = pynvml.nvmlDeviceGetHandleByIndex(index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (ncpu + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = sorted(c for c in cpus if c in allowed)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


def broadcast_bytes(dist, payload: bytes | None, n: int, src: int = 0) -> bytes:
    """Every rank gets rank `src`'s `n` bytes."""
    box = [payload if dist.get_rank() == src else None]
    dist.broadcast_object_list(box, src=src)
    out = bytes(box[0])
    if len(out) != n:
        raise FluidError(f"broadcast_bytes: expected {n} bytes, got {len(out)}")
    return out


def all_gather_bytes(dist, payload):
    """List over ranks of each rank's (picklable) payload."""
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, payload)
    return out


def attach(sim, dist, lib=None) -> None:
    """Collective over `dist`'s default group: give the slab handle its NCCL communicator and map the
    peers' gather sources.  `lib` defaults to the loaded libfluid_cuda (a stand-in is injected by tests)."""
    L = lib if lib is not None else sim._L
    rank, world = dist.get_rank(), dist.get_world_size()
    if world != sim.nranks or rank != sim.rank:
        raise FluidError(f"process group ({rank}/{world}) does not match the slab handle ({sim.rank}/{sim.nranks})")
    if world == 1:
        return
    path = nccl_library_path().encode()
    uid = None
    if rank == 0:
        buf = C.create_string_buffer(128)
        sim._ck(L.fluid_comm_unique_id(path, buf))
        uid = buf.raw
    uid = broadcast_bytes(dist, uid, 128)
    sim._ck(L.fluid_comm_init(sim._h, path, uid))
    # CUDA IPC: the advection gather sources of every rank, and every solver buffer + the arrival flags of
    # the two z-neighbours (sweeps then hand their halo planes over by peer stores, see include/fluid_cuda.h)
    gather_ids = [FIELDS[name] for name in GATHER_SOURCES + GATHER_PARTNERS]
    mine = {}
    for aid in range(IPC_COUNT):
        buf = C.create_string_buffer(64)
        sim._ck(L.fluid_ipc_export(sim._h, aid, buf))
        mine[aid] = buf.raw
    for r, handles in enumerate(all_gather_bytes(dist, mine)):
        if r == rank:
            continue
        wanted = range(IPC_COUNT) if abs(r - rank) == 1 else gather_ids + [IPC_SYNC]
        for aid in wanted:
            sim._ck(L.fluid_ipc_import(sim._h, r, aid, handles[aid]))
    dist.barrier()


def reduce_stats(dist, st: dict) -> dict:
    """Combine per-slab run() statistics (simulation.cpp:73-90): sums add, minima / maxima fold."""
    parts = all_gather_bytes(dist, st)
    out = {}
    for k in st:
        vals = [p[k] for p in parts]
        out[k] = sum(vals) if k.endswith("_sum") else (min(vals) if k.endswith("_min") else max(vals))
    return out


def gather_global_state(sim, dist, names=None) -> dict:
    """Global padded fields on rank 0, merged from every rank's owned planes (other ranks get partial
    sums they should ignore).  The owned planes are disjoint and everything else is all-zero bits, so an
    INTEGER sum is a bitwise merge (a float sum would turn -0.0 into +0.0)."""
    import numpy as np
    import torch
    st = sim.state()
    out = {}
    cuda = dist.get_backend() == "nccl"
    for n in (names or list(st)):
        t = torch.from_numpy(st[n].view(np.int32).copy())
        if cuda:
            t = t.cuda()
        dist.reduce(t, 0, op=dist.ReduceOp.SUM)
        out[n] = t.cpu().numpy().view(np.float32)
    return out
