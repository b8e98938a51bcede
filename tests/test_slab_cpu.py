"""CPU tests of the multi-rank host logic (world_size 2, gloo): slab partition and halo plan, the
rendezvous helpers of multi_gpu.py, `attach` against a recording stand-in for libfluid_cuda, and a numpy
emulation of the two-plane halo exchange driven by the same plan the CUDA library implements
(csrc/comm.cuh: fl_halo_exchange)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fluid_simulation_b200 import slab, multi_gpu  # noqa: E402


@pytest.mark.parametrize("d,n", [(256, 8), (17, 2), (9, 4), (512, 3), (16, 8)])
def test_partition_covers_depth(d, n):
    parts = slab.partition(d, n)
    assert parts[0][0] == 1 and sum(dl for _, dl in parts) == d
    for (z0, d0), (z1, _) in zip(parts, parts[1:]):
        assert z1 == z0 + d0
    assert max(dl for _, dl in parts) - min(dl for _, dl in parts) <= 1


def test_halo_plan_is_symmetric():
    d, n = 37, 4
    for r in range(n - 1):
        up, down = slab.halo_plan(d, n, r)["up"], slab.halo_plan(d, n, r + 1)["down"]
        assert up["peer"] == r + 1 and down["peer"] == r
        assert len(up["send_planes"]) == len(down["recv_planes"]) == slab.HALO == 4
        dl = slab.halo_plan(d, n, r)["d_local"]
        assert up["send_planes"] == (dl - 3, dl - 2, dl - 1, dl) and down["recv_planes"] == (-3, -2, -1, 0)
    assert slab.halo_plan(d, n, 0)["down"] is None and slab.halo_plan(d, n, n - 1)["up"] is None


class _FakeSim:
    """What multi_gpu.attach needs of a Simulation."""
    def __init__(self, rank, nranks):
        self.rank, self.nranks, self._h = rank, nranks, 1234 + rank
        self._L = None

    def _ck(self, rc):
        assert rc == 0


class _FakeLib:
    """Records the C-ABI calls attach() makes and hands out recognisable ids / handles."""
    def __init__(self, rank):
        self.rank, self.calls, self.imported = rank, [], {}

    def fluid_comm_unique_id(self, path, buf):
        C.memmove(buf, bytes([7]) * 128, 128); self.calls.append("uid"); return 0

    def fluid_comm_init(self, h, path, uid):
        self.calls.append(("init", bytes(uid))); return 0

    def fluid_ipc_export(self, h, fid, buf):
        C.memmove(buf, bytes([self.rank * 16 + fid]) * 64, 64); self.calls.append(("export", fid)); return 0

    def fluid_ipc_import(self, h, peer, fid, handle):
        self.imported[(peer, fid)] = bytes(handle); return 0


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. rendezvous helpers
        got = multi_gpu.broadcast_bytes(dist, bytes(range(128)) if rank == 0 else None, 128)
        assert got == bytes(range(128))
        parts = multi_gpu.all_gather_bytes(dist, {"r": rank})
        assert [p["r"] for p in parts] == list(range(world))
        # 2. attach against the recording library
        lib, sim = _FakeLib(rank), _FakeSim(rank, world)
        multi_gpu.attach(sim, dist, lib=lib)
        assert ("init", bytes([7]) * 128) in lib.calls
        assert ("uid" in lib.calls) == (rank == 0)
        from fluid_simulation_b200.simulation import FIELDS
        for peer in range(world):
            for name in multi_gpu.GATHER_SOURCES:
                key = (peer, FIELDS[name])
                if peer == rank:
                    assert key not in lib.imported
                else:
                    assert lib.imported[key] == bytes([peer * 16 + FIELDS[name]]) * 64
        # 3. four-plane halo exchange of a global array, numpy emulation of fl_halo_exchange
        d, hgt, w = 4 * world + 3, 4, 5
        glob = np.arange((d + 2) * hgt * w, dtype=np.float32).reshape(d + 2, hgt, w)
        plan = slab.halo_plan(d, world, rank)
        z0, dl = plan["z_lo"], plan["d_local"]
        HP = slab.HALO - 1
        local = np.full((dl + 2 * slab.HALO, hgt, w), -1.0, np.float32)   # local planes -3 .. dl+4 -> index +3
        local[HP + 1:dl + HP + 1] = glob[z0:z0 + dl]                    # owned interior planes only
        import torch
        reqs = []
        for side in ("down", "up"):
            p = plan[side]
            if p is None:
                continue
            send = torch.from_numpy(np.ascontiguousarray(local[p["send_planes"][0] + HP:p["send_planes"][-1] + HP + 1]))
            recv = torch.empty_like(send)
            reqs.append((dist.isend(send, p["peer"]), dist.irecv(recv, p["peer"]), recv, p))
        for s, r, recv, p in reqs:
            s.wait(); r.wait()
            local[p["recv_planes"][0] + HP:p["recv_planes"][-1] + HP + 1] = recv.numpy()
        for zl in range(-HP, dl + slab.HALO + 1):
            gz = z0 + zl - 1
            if 1 <= gz <= d and not (1 <= zl <= dl):                    # a halo plane interior to the tunnel
                assert np.array_equal(local[zl + HP], glob[gz]), (rank, zl)
        # 4. statistics fold
        st = multi_gpu.reduce_stats(dist, {"dens_sum": 1.0 + rank, "dens_min": -rank, "dens_max": rank})
        assert st == {"dens_sum": sum(1.0 + r for r in range(world)), "dens_min": -(world - 1), "dens_max": world - 1}
        q.put((rank, "ok"))
    except Exception as ex:  # report instead of hanging the peer
        import traceback
        q.put((rank, "FAIL " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_multi_rank_host_logic(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(msg == "ok" for _, msg in results), results
