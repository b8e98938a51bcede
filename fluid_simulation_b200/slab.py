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
    """Neighbour ranks and the LOCAL plane indices exchanged per halo update (two planes per side: one
    fused red-black sweep recomputes the colour-0 update of the neighbour's boundary plane from the second
    halo plane): send local planes 1,2 down / receive into planes -1,0; send planes d_local-1,d_local up /
    receive into d_local+1,d_local+2.  Global ghost planes (z=0, z=D+1) belong to the end slabs and are not
    exchanged.  The same rule is implemented by libfluid_cuda's fl_halo_exchange (csrc/comm.cuh)."""
    z_lo, dl = partition(depth, nranks)[rank]
    return {
        "z_lo": z_lo, "d_local": dl,
        "down": None if rank == 0 else {"peer": rank - 1, "send_planes": (1, 2), "recv_planes": (-1, 0)},
        "up": None if rank == nranks - 1 else {"peer": rank + 1, "send_planes": (dl - 1, dl),
                                               "recv_planes": (dl + 1, dl + 2)},
    }


def attach_nccl(sim, dist):
    """Give a slab handle its communicator (implemented with the multi-GPU path)."""
    from . import multi_gpu
    multi_gpu.attach(sim, dist)
