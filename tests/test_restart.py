"""Restart from a dumped frame: the five fields of a frame (simulation.cpp:140-148) are a
complete state, because step() overwrites the other six before reading them (simulation.cpp:96-150).
CPU: that property on the reference itself and on the restatement.  GPU: the Python and C++ hosts restart
from their own frame files and reproduce the uninterrupted run byte for byte; the meta.json sidecar."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle
from fluid_simulation_b200 import masks
from tests.util import assert_bitwise

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FRAME_FIELDS = ("dens", "obs", "v_x", "v_y", "v_z")


def _carry(make):
    w, h, d = 14, 10, 8
    a = make(w, h, d)
    a.set("obs", masks.sphere(w, h, d))
    a.run_iterations(3)
    frame = {f: a.get(f) for f in FRAME_FIELDS}
    a.run_iterations(2)
    b = make(w, h, d)
    for f in FRAME_FIELDS:
        b.set(f, frame[f])
    b.run_iterations(2)
    assert_bitwise(b.state(), a.state(), FRAME_FIELDS, "restart from a frame")


@pytest.mark.skipif(not oracle.have_ref(), reason="reference not built here")
def test_frame_is_a_complete_state_reference():
    _carry(lambda w, h, d: oracle.RefSim(w, h, d, acc=4, threads=1))


def test_frame_is_a_complete_state_oracle():
    _carry(lambda w, h, d: oracle.OracleSim(w, h, d, acc=4, order=oracle.ORDER_RB))


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,d", [(64, 16, 8), (128, 16, 8)])
def test_restart_on_the_streaming_path(tmp_path, w, h, d):
    """W >= 64: the default production path (TMA streaming sweeps, v <-> v_prev buffer swap with the first sweep
    reading its guess from the right-hand side, the 4-field diffusion launch with its lazily allocated fourth
    partner).  A restart begins from zeroed scratch / prev buffers while the uninterrupted run has stale,
    permuted ones; the frames must not depend on that.  The restart frame carries NON-ZERO edge and corner
    ghosts (the reference never writes them, so they persist), which pins the edge-copy path."""
    from fluid_simulation_b200 import Simulation
    rng = np.random.default_rng(4)
    a = Simulation(w, h, d, 5, acc=5)
    a.set("obs", masks.box(w, h, d))
    for f in ("dens", "v_x", "v_y", "v_z"):                     # edges / corners of the padded volume
        v = np.zeros(a.shape, np.float32)
        e = rng.uniform(0.5, 1.5, a.shape).astype(np.float32)
        edge = np.zeros(a.shape, bool)
        edge[0, 0, :] = edge[0, -1, :] = edge[-1, 0, :] = edge[-1, -1, :] = True
        edge[0, :, 0] = edge[0, :, -1] = edge[-1, :, 0] = edge[-1, :, -1] = True
        edge[:, 0, 0] = edge[:, 0, -1] = edge[:, -1, 0] = edge[:, -1, -1] = True
        v[edge] = e[edge]
        a.set(f, v)
    a.run(str(tmp_path / "a"), verbose=False)
    b = Simulation(w, h, d, 2, acc=5)
    assert b.restart_from_frames(str(tmp_path / "a"), 2) == 2
    b.run(str(tmp_path / "b"), verbose=False)
    for name in ("data.bin", "obs.bin", "v_x.bin", "v_y.bin", "v_z.bin"):
        full = np.fromfile(tmp_path / "a" / name, dtype=np.float32).reshape(5, -1)
        tail = np.fromfile(tmp_path / "b" / name, dtype=np.float32).reshape(2, -1)
        assert np.array_equal(full[3:].view(np.uint32), tail.view(np.uint32)), name
    assert np.any(np.fromfile(tmp_path / "b" / "v_x.bin", dtype=np.float32).reshape(2, *a.shape)[1, 0, 0, :] != 0)


@pytest.mark.gpu
def test_python_restart_and_sidecar(tmp_path):
    from fluid_simulation_b200 import Simulation, FluidError
    w, h, d = 40, 24, 20
    a = Simulation(w, h, d, 5, acc=5)
    a.set("obs", masks.sphere(w, h, d))
    a.run(str(tmp_path / "a"), verbose=False)
    meta = json.load(open(tmp_path / "a" / "meta.json"))
    assert (meta["width"], meta["height"], meta["depth"], meta["frames"]) == (w, h, d, 5)
    assert meta["padded"] == [d + 2, h + 2, w + 2] and meta["acc"] == 5
    npad = (w + 2) * (h + 2) * (d + 2)
    assert os.path.getsize(tmp_path / "a" / "data.bin") == 4 * npad * meta["frames"]
    b = Simulation(w, h, d, 2, acc=5)
    assert b.restart_from_frames(str(tmp_path / "a"), 2) == 2
    b.run(str(tmp_path / "b"), verbose=False)
    for name in ("data.bin", "obs.bin", "v_x.bin", "v_y.bin", "v_z.bin"):
        full = np.fromfile(tmp_path / "a" / name, dtype=np.float32).reshape(5, -1)
        tail = np.fromfile(tmp_path / "b" / name, dtype=np.float32).reshape(2, -1)
        assert np.array_equal(full[3:].view(np.uint32), tail.view(np.uint32)), name
    c = Simulation(w, h, d, 1, acc=5)
    assert c.restart_from_frames(str(tmp_path / "a")) == 4             # default: the last frame
    assert_bitwise(c.state(), a.state(), FRAME_FIELDS, "last frame")
    with pytest.raises(FluidError):
        c.restart_from_frames(str(tmp_path / "a"), 7)
    with pytest.raises(FluidError):
        Simulation(w + 1, h, d, 1).restart_from_frames(str(tmp_path / "a"))


@pytest.mark.gpu
def test_cpp_restart(tmp_path):
    exe = os.path.join(ROOT, "fluid_simulation_b200", "host", "fluid_tunnel")
    if not os.path.exists(exe):
        pytest.skip("fluid_tunnel not built")
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    common = [exe, "--w", "72", "--h", "20", "--d", "16", "--acc", "4", "--quiet"]        # W >= 64: streaming path
    r = subprocess.run(common + ["--iter", "4", "--obstacle", "sphere", "--out", str(tmp_path / "a")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    r = subprocess.run(common + ["--iter", "3", "--restart", f"{tmp_path / 'a'}:0", "--out", str(tmp_path / "b")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    for name in ("data.bin", "obs.bin", "v_x.bin", "v_y.bin", "v_z.bin"):
        full = np.fromfile(tmp_path / "a" / name, dtype=np.uint32).reshape(4, -1)
        tail = np.fromfile(tmp_path / "b" / name, dtype=np.uint32).reshape(3, -1)
        assert np.array_equal(full[1:], tail), name
    meta = json.load(open(tmp_path / "b" / "meta.json"))
    assert meta["frames"] == 3 and meta["width"] == 72
    r = subprocess.run(common + ["--iter", "1", "--restart", f"{tmp_path / 'a'}:9", "--out", str(tmp_path / "b")], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "frame out of range" in r.stderr
