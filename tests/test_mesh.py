"""generate_obstacle_mesh (GUI/utils.py:10-38): marching cubes of the padded obstacle volume at level 0.5.

skimage is not available, so nothing here is compared with the reference's own output (parity unpinned, see
oracle/mesh_oracle.c).  What the DATA determines is checked on the C restatement, and the CUDA kernels must
reproduce the restatement exactly:
  * the vertex set = the midpoints of all grid edges with one solid and one fluid end (computed with NumPy),
  * the surface is closed and consistently oriented (every directed edge is matched by its reverse),
  * normals point from solid to fluid ('descent'): the enclosed signed volume is positive and close to the solid
    volume; a sphere has Euler characteristic 2; obstacles without ambiguous cubes give a 2-manifold,
  * all 256 cube configurations, alone in an empty volume, give closed surfaces with the table's triangle count."""
import numpy as np
import pytest

from oracle import postproc as O


def crossing_midpoints(obs):
    """All grid-edge midpoints between a cell > 0.5 and a cell <= 0.5, as rows [x, y, z]."""
    s = obs > 0.5                                   # (nz, ny, nx)
    out = []
    for axis, (dz, dy, dx) in enumerate([(0, 0, 1), (0, 1, 0), (1, 0, 0)]):          # x, y, z edges
        a = s[: s.shape[0] - dz, : s.shape[1] - dy, : s.shape[2] - dx]
        b = s[dz:, dy:, dx:]
        z, y, x = np.nonzero(a != b)
        p = np.stack([x, y, z], axis=1).astype(np.float32)
        p[:, axis] += 0.5
        out.append(p)
    return np.concatenate(out)


def as_set(v):
    return set(map(tuple, np.asarray(v, dtype=np.float32).tolist()))


def directed_edges(faces):
    e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]]).astype(np.int64)
    return e


def assert_closed_and_oriented(faces, tag):
    e = directed_edges(faces)
    assert (e[:, 0] != e[:, 1]).all(), tag
    n = int(faces.max()) + 1
    fwd = np.sort(e[:, 0] * n + e[:, 1])
    rev = np.sort(e[:, 1] * n + e[:, 0])
    assert np.array_equal(fwd, rev), f"{tag}: a directed edge without its reverse (open or inconsistently oriented surface)"


def signed_volume(verts, faces):
    v = verts.astype(np.float64)
    a, b, c = v[faces[:, 0]], v[faces[:, 1]], v[faces[:, 2]]
    return float(np.einsum("ij,ij->i", a, np.cross(b, c)).sum() / 6.0)


def volumes():
    from fluid_simulation_b200 import masks
    rng = np.random.default_rng(5)
    rnd = np.zeros((14, 12, 16), np.float32)
    rnd[1:-1, 1:-1, 1:-1] = (rng.random((12, 10, 14)) < 0.35)
    box = np.zeros((9, 10, 11), np.float32)
    box[2:6, 3:8, 4:9] = 1.0
    return {"sphere": masks.sphere(40, 32, 24), "airfoil": masks.airfoil(48, 32, 16), "box": box, "random": rnd,
            "empty": np.zeros((6, 7, 8), np.float32)}


@pytest.mark.parametrize("name", ["sphere", "airfoil", "box", "random"])
def test_restatement_properties(name):
    obs = volumes()[name]
    verts, faces = O.marching_cubes(obs)
    want = crossing_midpoints(obs)
    assert len(verts) == len(want) and as_set(verts) == as_set(want)          # the vertex SET, and no duplicates
    assert faces.min() >= 0 and faces.max() < len(verts)
    assert len(np.unique(faces)) == len(verts)                                # every vertex is used
    assert_closed_and_oriented(faces, name)
    vol = signed_volume(verts, faces)
    solid = float((obs > 0.5).sum())
    assert vol > 0, f"{name}: normals point into the obstacle"
    if name != "random":
        assert 0.6 * solid <= vol <= solid, (vol, solid)
        e = directed_edges(faces)
        und = np.sort(e, axis=1)
        _, counts = np.unique(und, axis=0, return_counts=True)
        assert (counts == 2).all()                                            # 2-manifold
        if name in ("sphere", "box"):
            assert len(verts) - len(counts) + len(faces) == 2                 # Euler characteristic of a sphere


def test_every_cube_configuration_is_closed():
    ntri = []
    for case in range(256):
        obs = np.zeros((4, 4, 4), np.float32)
        for i in range(8):
            if (case >> i) & 1:
                obs[1 + ((i >> 2) & 1), 1 + ((i >> 1) & 1), 1 + (i & 1)] = 1.0
        verts, faces = O.marching_cubes(obs)
        assert as_set(verts) == as_set(crossing_midpoints(obs))
        if case:
            assert_closed_and_oriented(faces, f"case {case}")
            assert signed_volume(verts, faces) > 0
        else:
            assert len(verts) == 0 and len(faces) == 0
        ntri.append(len(faces))
    assert ntri[0] == 0 and ntri[1] == 8 and ntri[255] == 44       # one cell: an octahedron; a 2x2x2 block: 6 faces x 4 quads - 4 corners... counted once


def test_empty_volume_gives_the_empty_mesh():
    verts, faces = O.marching_cubes(volumes()["empty"])
    assert verts.shape == (0, 3) and faces.shape == (0, 3)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["sphere", "airfoil", "random", "empty"])
def test_gpu_mesh_equals_restatement(name):
    from fluid_simulation_b200 import Simulation
    obs = volumes()[name]
    d, h, w = (s - 2 for s in obs.shape)
    sim = Simulation(w, h, d, 1, acc=2)
    sim.set("obs", obs)
    mesh = sim.obstacle_mesh()
    verts, faces = O.marching_cubes(obs)
    if name == "empty":
        assert mesh["vertexes"].size == 0 and mesh["faces"].size == 0 and mesh["vertex_colors"].size == 0
        return
    assert np.array_equal(mesh["vertexes"], verts)          # same cell order, same interpolation: bit-identical
    assert np.array_equal(mesh["faces"], faces)
    assert mesh["vertex_colors"].shape == (len(verts), 4) and (mesh["vertex_colors"][:, :3] == 0.5).all()
    assert as_set(mesh["vertexes"]) == as_set(crossing_midpoints(obs))


@pytest.mark.gpu
def test_gpu_mesh_headline_size_and_buffer_errors():
    import ctypes as C
    from fluid_simulation_b200 import Simulation, masks, _lib
    w, h, d = 256, 256, 64
    obs = masks.airfoil(w, h, d)
    sim = Simulation(w, h, d, 1, acc=2)
    sim.set("obs", obs)
    mesh = sim.obstacle_mesh()
    assert as_set(mesh["vertexes"]) == as_set(crossing_midpoints(obs))
    assert_closed_and_oriented(mesh["faces"], "256x256x64 airfoil")
    L = _lib.load()
    nv, nf = C.c_longlong(0), C.c_longlong(0)
    buf = np.zeros(12, np.float32); ibuf = np.zeros(12, np.int32)
    rc = L.fluid_obstacle_mesh(sim._h, 0.5, 4, 4, C.byref(nv), C.byref(nf), buf.ctypes.data, ibuf.ctypes.data)
    assert rc == -1 and nv.value == len(mesh["vertexes"]) and nf.value == len(mesh["faces"])
    assert not buf.any() and not ibuf.any()                  # nothing copied
    assert b"do not fit" in L.fluid_last_error(sim._h)
