"""z-slab decomposition helpers for multi-GPU runs: the host-side plan of which
global planes each rank owns and which halo planes it exchanges with which neighbour."""
from __future__ import annotations


def partition(depth: int, nranks: int):
    """Equal z-slabs, remainder planes to the low ranks.  Returns [(z_lo, d_local)] with 1-based
    global z_lo -- the same rule libfluid_cuda's fluid_create_slab applies."""
    base, rem = divmod(depth, nranks)
    out, z = [], 1
    for r in range(nranks):
        dl = base + (1 if r < rem else 0)
        out.append((z, dl))
        z += dl
    return out


def halo_plan(depth: int, nranks: int, rank: int):
    """Neighbour ranks and the LOCAL plane indices exchanged per halo update:
    send local plane 1 down / receive into local plane 0, send plane d_local up / receive into
    plane d_local+1.  Global ghost planes (z=0, z=D+1) belong to the end slabs and are not exchanged."""
    z_lo, dl = partition(depth, nranks)[rank]
    return {
        "z_lo": z_lo, "d_local": dl,
        "down": None if rank == 0 else {"peer": rank - 1, "send_plane": 1, "recv_plane": 0},
        "up": None if rank == nranks - 1 else {"peer": rank + 1, "send_plane": dl, "recv_plane": dl + 1},
    }


def attach_nccl(sim, dist):
    """Give a slab handle its communicator (implemented with the multi-GPU path)."""
    from . import multi_gpu
    multi_gpu.attach(sim, dist)
