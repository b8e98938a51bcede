"""The temporally blocked sweep kernel (two red-black sweeps per HBM pass, csrc/sweep_pair_kernel.cuh) compiled
for the HOST (tests/host_emul/pair_emul.cpp: one host thread per CUDA thread, barriers / shuffles / TMA
emulated) against the ordering-matched CPU oracle, bit for bit: linearSolver with two iterations
(simulation.cpp:251-273) for every setBounds flavour, both streaming directions, z-chunks that cut the
columns into short segments, masks with solids on the walls, several row tiles."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import OracleSim, ORDER_RB

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def emul():
    out = os.path.join(HERE, "host_emul", "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libpair_emul.so")
    src = os.path.join(HERE, "host_emul", "pair_emul.cpp")
    deps = [src] + [os.path.join(HERE, "..", "fluid_simulation_b200", "csrc", f)
                    for f in ("sweep_pair_kernel.cuh", "sweep_common.cuh", "common.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.run(["g++", "-std=c++20", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-pthread",
                        "-I/usr/local/cuda/include", "-Wno-attributes", src, "-o", so], check=True)
    L = C.CDLL(so)
    L.pair_emul_sweep.argtypes = [C.c_int] * 3 + [C.c_void_p] * 3 + [C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p]
    return L


def run_case(L, w, h, d, b, zchunk, up, seed=5, wall_solids=False, unroll=6):
    from fluid_simulation_b200 import masks
    st = masks.random_state(w, h, d, seed=seed)
    if wall_solids:                                    # solids touching every wall: flagged cells next to ghosts
        obs = st["obs"]
        rng = np.random.default_rng(seed + 1)
        obs[1:-1, 1:-1, 1:-1] = (rng.uniform(size=(d, h, w)) < 0.15).astype(np.float32)
        st["obs"] = obs
    cpu = OracleSim(w, h, d, acc=2, order=ORDER_RB)
    cpu.load_state(st)
    cpu.linear_solver(b, "v_y", "v_y_prev", 0.37, 3.22)
    want = cpu.get("v_y")
    x = np.ascontiguousarray(st["v_y"]); rhs = np.ascontiguousarray(st["v_y_prev"]); obs = np.ascontiguousarray(st["obs"])
    out = np.zeros_like(x)
    rc = L.pair_emul_sweep(w, h, d, x.ctypes.data, rhs.ctypes.data, obs.ctypes.data, b, 0.37, 3.22, zchunk, up, unroll, out.ctypes.data)
    assert rc == 0
    bad = out.view(np.uint32) != want.view(np.uint32)
    if bad.any():
        idx = np.argwhere(bad)
        zs, ys, xs = (sorted(set(idx[:, k].tolist())) for k in range(3))
        raise AssertionError(f"{w}x{h}x{d} b={b} zchunk={zchunk} up={up}: {bad.sum()} cells differ; first (z,y,x)={tuple(idx[0])} "
                             f"got {out[tuple(idx[0])]} want {want[tuple(idx[0])]}; z in {zs[:12]}, y in {ys[:12]}, x in {xs[:12]}")


@pytest.mark.parametrize("b", [0, 1, 2, 3])
@pytest.mark.parametrize("up", [1, 0])
def test_pair_sweep_small(emul, b, up):
    run_case(emul, 136, 12, 7, b, 7, up, wall_solids=True)


@pytest.mark.parametrize("w,h,d,zchunk", [(256, 24, 9, 4), (200, 10, 5, 1), (144, 2, 3, 3)])
def test_pair_sweep_shapes(emul, w, h, d, zchunk):
    run_case(emul, w, h, d, 2, zchunk, 1, wall_solids=True)
    run_case(emul, w, h, d, 3, zchunk, 0, unroll=2)
