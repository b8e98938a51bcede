"""The C++ host side of the drop-in (fluid_simulation_b200/host): the reference's Simulation API over
the C ABI, and the fluid_tunnel driver.  CPU part: it builds, exports the reference's class, the
reference's own STL loader compiles against its header, and without a GPU it fails loudly.  GPU part:
the frames it writes are byte-identical to what the CPU oracle computes, in the reference's file format."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "fluid_simulation_b200", "host")
TUNNEL = os.path.join(HOST, "fluid_tunnel")


def test_host_library_exports_reference_class():
    out = subprocess.run(["nm", "-DC", os.path.join(HOST, "libfluid_host.so")], capture_output=True, text=True).stdout
    for sym in ("Simulation::Simulation(int, int, int, int, int, float, float, float, int, unsigned int)",
                "Simulation::run()", "Simulation::step()", "Simulation::addObstacle(int, int, int)",
                "Simulation::addDensity(int, int, int, float)",
                "Simulation::setVelocity(int, int, int, float, float, float)"):
        assert sym in out, sym


@pytest.mark.skipif(not os.path.exists("/root/reference/object_loader.cpp"), reason="reference tree absent")
def test_reference_loader_compiles_against_our_header(tmp_path):
    """object_loader.cpp is used UNMODIFIED, where it lies: it only needs width/height/depth + addObstacle."""
    res = subprocess.run(["c++", "-std=c++20", "-O1", "-fopenmp", "-I", HOST, "-c", "/root/reference/object_loader.cpp",
                          "-o", str(tmp_path / "loader.o")], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-2000:]


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="checks the no-GPU failure mode")
def test_driver_fails_loudly_without_gpu(tmp_path):
    res = subprocess.run([TUNNEL, "--w", "8", "--h", "8", "--d", "8", "--iter", "1", "--out", str(tmp_path)],
                         capture_output=True, text=True)
    assert res.returncode == 1
    assert "no CPU fallback" in res.stderr or "CUDA" in res.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("solver,order_name", [("rb", "ORDER_RB"), ("lex", "ORDER_LEX")])
def test_driver_frames_match_oracle(tmp_path, solver, order_name):
    import oracle
    from fluid_simulation_b200 import masks
    w, h, d, acc, iters = 24, 16, 12, 5, 4
    out = tmp_path / "data"; out.mkdir()
    res = subprocess.run([TUNNEL, "--w", str(w), "--h", str(h), "--d", str(d), "--iter", str(iters), "--acc", str(acc),
                          "--obstacle", "sphere", "--solver", solver, "--out", str(out)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    assert "simulation finished" in res.stdout
    cpu = oracle.OracleSim(w, h, d, acc=acc, order=getattr(oracle, order_name))
    cpu.set("obs", masks.sphere(w, h, d))
    frames = {n: [] for n in ("dens", "obs", "v_x", "v_y", "v_z")}
    for _ in range(iters):
        cpu.run_iterations(1)
        for n in frames:
            frames[n].append(cpu.get(n).copy())
    size = (w + 2) * (h + 2) * (d + 2)
    for fname, n in (("data.bin", "dens"), ("obs.bin", "obs"), ("v_x.bin", "v_x"), ("v_y.bin", "v_y"), ("v_z.bin", "v_z")):
        got = np.fromfile(out / fname, dtype=np.float32)
        assert got.size == size * iters, (fname, got.size)          # gui.py:229 makes the same check
        want = np.concatenate([f.ravel() for f in frames[n]])
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), fname


def test_slab_hub_collectives(tmp_path):
    """host/slab_hub.h (the rendezvous of `fluid_tunnel --gpus N`, no CUDA involved): N forked ranks, an all-gather
    of rank-dependent payloads, a barrier and the parent's exit status = the worst rank's."""
    src = tmp_path / "hub.cpp"
    src.write_text('''#include "slab_hub.h"
#include <cstdio>
int main(int argc, char** argv) {
    const int n = 3;
    slab::Hub* hub = slab::spawn(n);                 // the parent serves and exits inside
    int mine[2] = { hub->rank * 10, hub->rank * 10 + 1 }, all[6] = {};
    hub->allgather(mine, sizeof mine, all);
    hub->barrier();
    for (int r = 0; r < n; ++r) if (all[2 * r] != r * 10 || all[2 * r + 1] != r * 10 + 1) return 9;
    double s = hub->rank, sums[3]; hub->allgather(&s, sizeof s, sums);
    if (sums[0] + sums[1] + sums[2] != 3.0) return 8;
    return argc > 1 && hub->rank == 2 ? 7 : 0;       // with an argument rank 2 fails: the parent must report it
}
''')
    exe = tmp_path / "hub"
    res = subprocess.run(["g++", "-std=c++17", "-O1", "-I", HOST, str(src), "-o", str(exe)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-2000:]
    assert subprocess.run([str(exe)], timeout=60).returncode == 0
    assert subprocess.run([str(exe), "fail"], timeout=60).returncode == 7


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("n", [2, 4])
def test_driver_multi_gpu_frames_equal_single_gpu(tmp_path, n):
    """fluid_tunnel --gpus N (C++ host, one forked process per GPU, no Python): the frame files, meta.json and the
    printed statistics equal the single-GPU run's byte for byte."""
    if _ngpu() < n:
        pytest.skip(f"needs {n} GPUs")
    env = dict(os.environ)
    try:
        import nvidia.nccl as nn
        env["FLUID_NCCL_LIB"] = os.path.join(os.path.dirname(nn.__file__), "lib", "libnccl.so.2")
    except Exception:
        pass
    common = [TUNNEL, "--w", "136", "--h", "20", "--d", str(36 * n), "--iter", "3", "--acc", "4", "--obstacle", "sphere"]
    outs = {}
    for tag, extra in (("one", []), ("many", ["--gpus", str(n)])):
        out = tmp_path / tag; out.mkdir()
        res = subprocess.run(common + extra + ["--out", str(out)], capture_output=True, text=True, timeout=600, env=env)
        assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
        outs[tag] = (out, res.stdout)
    for name in ("data.bin", "obs.bin", "v_x.bin", "v_y.bin", "v_z.bin"):
        a = np.fromfile(outs["one"][0] / name, dtype=np.uint32); b = np.fromfile(outs["many"][0] / name, dtype=np.uint32)
        assert a.size == b.size and np.array_equal(a, b), name
    stats = lambda s: [l for l in s.splitlines() if " min = " in l or " max = " in l]
    assert stats(outs["one"][1]) == stats(outs["many"][1]) and len(stats(outs["one"][1])) == 8
