"""STL voxeliser (object_loader.cpp:98-174 reader, :288-449 voxeliser).

CPU: the STL reader's acceptance rules; the plain-C restatement (oracle/voxel_oracle.c) against its committed
masks and, statistically, against the reference's own committed mask (the reference seeds jitter and rays
from the thread id, so bit parity with it does not exist); live against the reference where it is built.
GPU: the CUDA voxeliser against the restatement, bit for bit (mask and "added" count), in all three
sampling modes, through the Python mirror, the STL file path and the C++ host loader."""
import glob
import os
import subprocess

import numpy as np
import pytest

import oracle
from fluid_simulation_b200 import stl

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "vox_*.npz")))


def unpack(bits, w, h, d):
    n = (d + 2) * (h + 2) * (w + 2)
    return np.unpackbits(bits)[:n].reshape(d + 2, h + 2, w + 2).astype(bool)


def iou(a, b):
    u = (a | b).sum()
    return float((a & b).sum()) / float(u) if u else 1.0


def case(path):
    g = np.load(path)
    w, h, d = (int(v) for v in g["dims"])
    return g, w, h, d, float(g["scale"]), tuple(float(v) for v in g["rot"]), tuple(float(v) for v in g["translate"])


# ---- reader ------------------------------------------------------------------------------------
def test_reader_binary_and_ascii_roundtrip(tmp_path):
    tris = stl.sphere_mesh(1.5, 6, 10, center=(0.25, -1.0, 2.0))
    pb, pa = tmp_path / "m.stl", tmp_path / "a.stl"
    stl.write_stl_binary(pb, tris); stl.write_stl_ascii(pa, tris)
    assert np.array_equal(stl.read_stl(pb), tris)
    assert np.array_equal(stl.read_stl(pa), tris)


def test_reader_rules(tmp_path):
    # object_loader.cpp:107-110: "solid" at the start of the first line means ASCII, anything else binary
    p = tmp_path / "q.stl"
    p.write_text("  solid x\nfacet normal 0 0 1\nouter loop\nvertex 0 0 0\nvertex 1 0 0\nvertex 0 1 0\nendloop\nendfacet\n"
                 # four vertices: the counter wraps to 0 (:166) and the facet is dropped (:154)
                 "facet normal 0 0 1\nouter loop\nvertex 0 0 0\nvertex 1 0 0\nvertex 0 1 0\nvertex 9 9 9\nendloop\nendfacet\n"
                 # two vertices: dropped as well
                 "facet normal 0 0 1\nouter loop\nvertex 0 0 0\nvertex 1 0 0\nendloop\nendfacet\nendsolid x\n")
    t = stl.read_stl(p)
    assert t.shape == (1, 9) and np.array_equal(t[0], np.array([0, 0, 0, 1, 0, 0, 0, 1, 0], np.float32))
    # truncated binary file: what is complete is returned; shorter than a header: nothing
    tris = stl.box_mesh()
    pb = tmp_path / "b.stl"; stl.write_stl_binary(pb, tris)
    raw = pb.read_bytes(); (tmp_path / "t.stl").write_bytes(raw[:84 + 50 * 5 + 17])
    assert np.array_equal(stl.read_stl(tmp_path / "t.stl"), tris[:5])
    (tmp_path / "e.stl").write_bytes(b"xx")
    assert stl.read_stl(tmp_path / "e.stl").shape == (0, 9)


def test_meshes_are_closed():
    for tris in (stl.box_mesh(), stl.sphere_mesh(1.0, 8, 16)):
        e = {}
        for t in tris.reshape(-1, 3, 3):
            for a, b in ((0, 1), (1, 2), (2, 0)):
                k = tuple(sorted((tuple(np.round(t[a], 5)), tuple(np.round(t[b], 5)))))
                e[k] = e.get(k, 0) + 1
        assert set(e.values()) == {2}


# ---- oracle ------------------------------------------------------------------------------------
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-4] for p in GOLDEN])
def test_oracle_against_committed_masks(path):
    g, w, h, d, scale, rot, tr = case(path)
    for mode in (0, 1, 2):
        m, added, setup = oracle.voxelize(g["tris"], w, h, d, scale, rot, tr, mode)
        assert added == int(g[f"added{mode}"])
        assert np.array_equal(np.packbits(m > 0), g[f"mask{mode}"]), f"mode {mode}"
    assert [setup["nx"], setup["ny"], setup["nz"]] == [int(v) for v in g["lattice"]]
    assert np.float32(setup["resolution"]) == g["resolution"]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-4] for p in GOLDEN])
def test_oracle_matches_reference_mask_statistically(path):
    """Same lattice, same tests, different random jitter / rays: the masks differ only in a thin shell."""
    g, w, h, d, scale, rot, tr = case(path)
    ref = unpack(g["ref_mask"], w, h, d)
    for mode in (0, 1):
        mine = unpack(g[f"mask{mode}"], w, h, d)
        assert iou(ref, mine) >= 0.97, (mode, iou(ref, mine))
        assert abs(int(mine.sum()) - int(ref.sum())) <= 0.02 * ref.sum() + 2
    centre = unpack(g["mask2"], w, h, d)                       # centre rule: a subset, up to boundary cells
    assert (centre & ~ref).sum() <= 0.02 * ref.sum() + 2


@pytest.mark.skipif(not oracle.have_ref(), reason="reference not built here")
def test_oracle_matches_live_reference(tmp_path):
    tris = stl.box_mesh((-0.8, -0.6, -0.3), (0.9, 0.4, 0.5))
    p = tmp_path / "box.stl"; stl.write_stl_ascii(p, tris)
    ref = oracle.ref_voxelize(str(p), 44, 36, 30, 0.75, (10.0, 0.0, 25.0), (1.0, 2.0, -1.0)) > 0
    mine, added, _ = oracle.voxelize(stl.read_stl(p), 44, 36, 30, 0.75, (10.0, 0.0, 25.0), (1.0, 2.0, -1.0), 0)
    assert added > 0 and iou(ref, mine > 0) >= 0.97


def test_oracle_rejects_empty_mesh():
    _, added, _ = oracle.voxelize(np.zeros((0, 9), np.float32), 8, 8, 8)
    assert added == -1


# ---- CUDA --------------------------------------------------------------------------------------
def gpu_mask(w, h, d, tris, scale, rot, tr, mode):
    from fluid_simulation_b200 import Simulation
    sim = Simulation(w, h, d, 1)
    added = sim.voxelize(tris, scale, rot, tr, mode)
    return sim.get("obs"), added


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-4] for p in GOLDEN])
def test_cuda_voxeliser_matches_committed_masks(path, mode):
    g, w, h, d, scale, rot, tr = case(path)
    obs, added = gpu_mask(w, h, d, g["tris"], scale, rot, tr, mode)
    assert added == int(g[f"added{mode}"])
    assert np.array_equal(np.packbits(obs > 0), g[f"mask{mode}"])
    assert set(np.unique(obs)) <= {0.0, 1.0}


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("dims,scale,rot,tr", [
    ((64, 48, 40), 0.8, (0.0, 0.0, 0.0), (0.0, 0.0, 0.0)),
    ((52, 60, 44), 0.9, (33.0, -20.0, 75.0), (-4.0, 6.5, 2.0)),
    ((30, 30, 30), 1.6, (0.0, 45.0, 0.0), (10.0, 0.0, -9.0)),          # partly outside the grid
    ((33, 17, 9), 0.5, (90.0, 0.0, 0.0), (0.0, 0.0, 0.0)),
])
def test_cuda_voxeliser_matches_oracle(dims, scale, rot, tr, mode):
    w, h, d = dims
    tris = stl.sphere_mesh(1.0, 7, 12, center=(0.2, -0.1, 0.3)) if mode != 1 else stl.box_mesh()
    want, want_added, _ = oracle.voxelize(tris, w, h, d, scale, rot, tr, mode)
    obs, added = gpu_mask(w, h, d, tris, scale, rot, tr, mode)
    assert added == want_added
    assert np.array_equal(obs, want)


@pytest.mark.gpu
def test_cuda_voxeliser_many_triangles_and_step():
    """More triangles than one shared-memory tile; the mask then drives the solver like any other obstacle."""
    from fluid_simulation_b200 import Simulation
    w, h, d = 40, 40, 40
    tris = stl.sphere_mesh(1.0, 20, 40)                         # 1520 triangles
    want, want_added, _ = oracle.voxelize(tris, w, h, d, 0.8, (0, 0, 0), (0, 0, 0), 0)
    sim = Simulation(w, h, d, 1, acc=4)
    assert sim.voxelize(tris, 0.8) == want_added
    assert np.array_equal(sim.get("obs"), want)
    cpu = oracle.OracleSim(w, h, d, acc=4, order=oracle.ORDER_RB)
    cpu.set("obs", want)
    sim.run_iterations(2); cpu.run_iterations(2)
    from tests.util import assert_bitwise
    assert_bitwise(sim.state(), cpu.state(), oracle.FIELD_NAMES, "2 iterations around a voxelised sphere")


@pytest.mark.gpu
def test_load_stl_paths(tmp_path, capsys):
    from fluid_simulation_b200 import Simulation
    tris = stl.box_mesh()
    p = tmp_path / "b.stl"; stl.write_stl_binary(p, tris)
    sim = Simulation(32, 24, 20, 1)
    added = sim.load_stl(str(p), 0.8, (0.0, 30.0, 0.0), (1.0, 0.0, 0.0))
    want, want_added, _ = oracle.voxelize(tris, 32, 24, 20, 0.8, (0.0, 30.0, 0.0), (1.0, 0.0, 0.0), 0)
    assert added == want_added and np.array_equal(sim.get("obs"), want)
    # unreadable / empty files are reported and skipped (object_loader.cpp:292-296); the mask is untouched
    assert sim.load_stl(str(tmp_path / "missing.stl")) == 0
    assert "Failed to load STL" in capsys.readouterr().err
    assert np.array_equal(sim.get("obs"), want)
    with pytest.raises(Exception):
        sim.voxelize(np.zeros((0, 9), np.float32))


@pytest.mark.gpu
def test_cpp_loader_matches_oracle(tmp_path):
    """host/object_loader.cpp (the reference's loadSTLIntoObstacles signature) through fluid_tunnel."""
    exe = os.path.join(ROOT, "fluid_simulation_b200", "host", "fluid_tunnel")
    if not os.path.exists(exe):
        pytest.skip("fluid_tunnel not built")
    tris = stl.sphere_mesh(1.0, 8, 14)
    p = tmp_path / "s.stl"; stl.write_stl_ascii(p, tris)
    out = tmp_path / "data"; out.mkdir()
    w, h, d = 36, 28, 24
    r = subprocess.run([exe, "--w", str(w), "--h", str(h), "--d", str(d), "--iter", "1", "--acc", "2", "--obstacle", str(p),
                        "--stl-scale", "0.7", "--stl-rot", "15,0,40", "--stl-translate", "2,-1,0", "--out", str(out), "--quiet"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    got = np.fromfile(out / "obs.bin", dtype=np.float32).reshape(-1, d + 2, h + 2, w + 2)[0]
    want, _, _ = oracle.voxelize(stl.read_stl(p), w, h, d, 0.7, (15.0, 0.0, 40.0), (2.0, -1.0, 0.0), 0)
    assert np.array_equal(got, want)
