"""z-slab decomposition helpers for multi-GPU runs: the host-side plan of which
global planes each rank owns and which halo planes it exchanges with which neighbour."""
from __future__ import annotations


HALO = 4        # FL_HALO of csrc/common.cuh


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
    """Neighbour ranks and the LOCAL plane indices exchanged per halo update.  A slab keeps HALO = 4 planes per
    interior side (one fused red-black sweep needs two: the colour-0 update of the neighbour's boundary plane is
    recomputed from the second halo plane; the pass that fuses two sweeps needs four): send local planes
    1..4 down / receive into planes -3..0; send planes d_local-3..d_local up / receive into
    d_local+1..d_local+4.  Global ghost planes (z=0, z=D+1) belong to the end slabs and are not exchanged.
    The same rule is implemented by libfluid_cuda's fl_halo_exchange (csrc/comm.cuh); every slab has at least
    HALO planes."""
    z_lo, dl = partition(depth, nranks)[rank]
    return {
        "z_lo": z_lo, "d_local": dl,
        "down": None if rank == 0 else {"peer": rank - 1, "send_planes": tuple(range(1, HALO + 1)),
                                        "recv_planes": tuple(range(1 - HALO, 1))},
        "up": None if rank == nranks - 1 else {"peer": rank + 1, "send_planes": tuple(range(dl - HALO + 1, dl + 1)),
                                               "recv_planes": tuple(range(dl + 1, dl + HALO + 1))},
    }


def attach_nccl(sim, dist):
    """Give a slab handle its communicator (implemented with the multi-GPU path)."""
    from . import multi_gpu
    multi_gpu.attach(sim, dist)
