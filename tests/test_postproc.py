"""The offline viewers' arithmetic: streamlines (GUI/utils.py:40-214) and the colour-mapped
z-slice (gui.py:61-79, 265-300).

CPU: the C restatement (oracle/postproc_oracle.c) against the committed output of the reference's own
GUI/utils.py (tests/golden/pp_*.npz; tolerance 1e-9 on positions because np.linalg.norm's BLAS summation is
unspecified, the accepted lines and their lengths must be identical) and live against the reference where it
is present; the product's per-seed source (csrc/postproc_core.h) compiled for the host, bit for bit against the
restatement; the two independent restatements of matplotlib's colour table against each other.
GPU: the CUDA kernels through the C ABI, bit for bit against the restatement."""
import ctypes as C
import glob
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import postproc as opp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "pp_*.npz")))
IDS = [os.path.basename(p)[3:-4] for p in GOLDEN]
TOL = 1e-9          # positions / colour argument vs the reference's Python (see module docstring)


def load(path):
    g = np.load(path)
    d, prox, steps, ss, thr = g["params"]
    return g, dict(density=int(d), proximity=float(prox), integration_steps=int(steps), step_size=float(ss)), float(thr)


def oracle_lines(g, kw, thr):
    return opp.streamlines(g["vx"], g["vy"], g["vz"], g["obs"], threshold=thr, **kw)


def same_lines(a, b, tol=0.0):
    (la, na), (lb, nb) = a, b
    assert [len(x) for x in la] == [len(x) for x in lb]
    if not la:
        return
    pa, pb = np.concatenate(la), np.concatenate(lb)
    if tol == 0.0:
        assert np.array_equal(pa.view(np.uint64), pb.view(np.uint64))
        assert np.array_equal(np.asarray(na).view(np.uint64), np.asarray(nb).view(np.uint64))
    else:
        assert np.abs(pa - pb).max() <= tol and np.abs(np.asarray(na) - np.asarray(nb)).max() <= tol


# ---- checker vs the reference ---------------------------------------------------------------------------
@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_oracle_against_reference_golden(path):
    g, kw, thr = load(path)
    assert len(g["line_len"]) > 0
    lines, norm = oracle_lines(g, kw, thr)
    same_lines((lines, norm), (np.split(g["points"], np.cumsum(g["line_len"])[:-1]), g["norm_speed"]), TOL)


@pytest.mark.skipif(not os.path.exists("/root/reference/GUI/utils.py"), reason="reference absent (GPU box)")
def test_oracle_against_reference_live():
    sys.path.insert(0, os.path.join(HERE, "golden"))
    try:
        import make_postproc_golden as mk
    finally:
        sys.path.pop(0)
    vx, vy, vz, obs = mk.fields(30, 22, 18, seed=21)
    ref_lines, ref_norm, pos, samp = mk.run_reference(vx, vy, vz, obs, 14, 3, 50, 0.25, 0.05)
    assert len(ref_lines) > 0
    same_lines(opp.streamlines(vx, vy, vz, obs, 14, 3, 50, 0.25, 0.05), (ref_lines, ref_norm), TOL)


def test_oracle_edge_cases():
    g, kw, thr = load(GOLDEN[0])
    zero = np.zeros_like(g["obs"])
    assert opp.streamlines(g["vx"], g["vy"], g["vz"], zero, threshold=thr, **kw)[0] == []          # no obstacle: no lines
    assert opp.streamlines(zero, zero, zero, g["obs"], threshold=thr, **kw)[0] == []               # fluid at rest
    assert opp.streamlines(g["vx"], g["vy"], g["vz"], g["obs"], 1, 2, 60, 0.2, thr)[0] == []        # density // 2 == 0


# ---- the product's per-seed source on the host ----------------------------------------------------------
@pytest.fixture(scope="module")
def emul():
    out = os.path.join(HERE, "host_emul", "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libpp_emul.so")
    subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC",
                    os.path.join(HERE, "host_emul", "pp_emul.cpp"), "-o", so], check=True)
    L = C.CDLL(so)
    L.emul_streamlines.restype = C.c_int
    L.emul_streamlines.argtypes = [C.c_void_p] * 4 + [C.c_int] * 5 + [C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_double,
                                                                       C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_longlong)]
    return L


def emul_lines(L, g, kw, thr):
    vols = [np.ascontiguousarray(g[k], np.float32) for k in ("vx", "vy", "vz", "obs")]
    nz, ny, nx = vols[0].shape
    zz, yy, xx = np.nonzero(vols[3] > 0.5)
    widen = kw["proximity"] / 10
    bmin = np.array([xx.min(), yy.min(), zz.min()], np.float64) - widen
    bmax = np.array([xx.max(), yy.max(), zz.max()], np.float64) + widen
    denom = float(np.float32(max(v.max() for v in vols[:3])) + np.float32(1e-6))
    d, steps = kw["density"], kw["integration_steps"]
    cap = max(1, d * (d // 2) ** 2)
    lens = np.zeros(cap, np.int32); norm = np.zeros(cap); pts = np.zeros((cap * (2 * (steps // 2) + 1), 3)); npts = C.c_longlong()
    n = L.emul_streamlines(*[v.ctypes.data for v in vols], nx, ny, nz, d, steps, kw["step_size"], thr, bmin.ctypes.data,
                           bmax.ctypes.data, denom, lens.ctypes.data, norm.ctypes.data, pts.ctypes.data, C.byref(npts))
    return (np.split(pts[:npts.value], np.cumsum(lens[:n])[:-1]) if n else []), norm[:n]


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_kernel_source_on_host_matches_oracle_bitwise(emul, path):
    g, kw, thr = load(path)
    same_lines(emul_lines(emul, g, kw, thr), oracle_lines(g, kw, thr))
    kw2 = dict(kw, integration_steps=7, density=kw["density"] + 3)          # odd step count, other lattice
    same_lines(emul_lines(emul, g, kw2, 0.0), oracle_lines(g, kw2, 0.0))


# ---- colour table and slice image -----------------------------------------------------------------------
def test_colour_table_restatements_agree():
    from fluid_simulation_b200 import viewer
    a, b = viewer.density_cmap_lut(), opp.density_cmap_lut()
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
    # anchors: white at 0, red at the top, pure blue at 4/6 of the range (index 170 = 255 * 4 / 6)
    assert np.array_equal(a[0], [1, 1, 1]) and np.array_equal(a[255], [1, 0, 0]) and np.allclose(a[170], [0, 0, 1])


def test_oracle_slice_basics():
    f = np.zeros((3, 4, 5), np.float32); obs = np.zeros_like(f)
    f[1, 0, :] = [-1.0, 0.0, 0.005, 0.01, 5.0]; f[1, 1, 0] = np.nan; obs[1, 2, :] = 1.0
    img = opp.render_slice(f, obs, 1, 0.0, 0.01)
    lut8 = (opp.density_cmap_lut() * 255).astype(np.uint8)
    assert np.array_equal(img[0, 0], lut8[0]) and np.array_equal(img[0, 1], lut8[0])            # clipped below, vmin
    assert np.array_equal(img[0, 2], lut8[127]) and np.array_equal(img[0, 3], lut8[255])        # float32(0.005) < 0.005; vmax -> N - 1
    assert np.array_equal(img[0, 4], lut8[255]) and np.array_equal(img[1, 0], [0, 0, 0])        # clipped above, NaN
    assert np.array_equal(img[2, 0], (lut8[0].astype(np.float32) * np.float32(0.8)).astype(np.uint8))   # darkened obstacle


# ---- GPU: the CUDA path through the C ABI ---------------------------------------------------------------
def gpu_sim(g):
    from fluid_simulation_b200 import Simulation
    nz, ny, nx = g["vx"].shape
    sim = Simulation(nx - 2, ny - 2, nz - 2, 0)
    for name, key in (("v_x", "vx"), ("v_y", "vy"), ("v_z", "vz"), ("obs", "obs")):
        sim.set(name, g[key])
    return sim


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_gpu_streamlines_bitwise(path):
    g, kw, thr = load(path)
    with gpu_sim(g) as sim:
        n0 = sim.launch_count()
        same_lines(sim.streamlines(velocity_change_threshold=thr, **kw), oracle_lines(g, kw, thr))
        assert sim.launch_count() > n0
        kw2 = dict(kw, integration_steps=7, density=kw["density"] + 3)
        same_lines(sim.streamlines(velocity_change_threshold=0.0, **kw2), oracle_lines(g, kw2, 0.0))
        # also against the reference's own output
        lines, norm = sim.streamlines(velocity_change_threshold=thr, **kw)
        same_lines((lines, norm), (np.split(g["points"], np.cumsum(g["line_len"])[:-1]), g["norm_speed"]), TOL)


@pytest.mark.gpu
def test_gpu_streamlines_edge_cases():
    g, kw, thr = load(GOLDEN[0])
    with gpu_sim(g) as sim:
        assert sim.streamlines(density=1)[0] == []                      # density // 2 == 0 seeds
        sim.set("obs", np.zeros_like(g["obs"]))
        assert sim.streamlines(**kw)[0] == []                           # no obstacle


@pytest.mark.gpu
def test_gpu_streamlines_after_a_run():
    """On a real solver state (not a synthetic field): 40 x 24 x 24 tunnel with a sphere, 6 steps."""
    from fluid_simulation_b200 import Simulation, masks
    with Simulation(40, 24, 24, 6, acc=10) as sim:
        sim.set("obs", masks.sphere(40, 24, 24))
        sim.run_iterations(6)
        st = {k: sim.get(k) for k in ("v_x", "v_y", "v_z", "obs", "dens")}
        kw = dict(density=16, proximity=20, integration_steps=60, step_size=0.2)
        got = sim.streamlines(velocity_change_threshold=0.01, **kw)
        want = opp.streamlines(st["v_x"], st["v_y"], st["v_z"], st["obs"], threshold=0.01, **kw)
        same_lines(got, want)
        for field, rng in (("dens", (0.0, 0.01)), ("v_x", (-10.0, 10.0)), ("v_y", (-1.0, 1.0))):
            for z in (0, 12, 25):
                img = sim.render_slice(field, z)
                assert np.array_equal(img, opp.render_slice(st[field], st["obs"], z, *rng))


@pytest.mark.gpu
def test_gpu_slice_special_values(tmp_path):
    from fluid_simulation_b200 import Simulation, viewer
    rng = np.random.default_rng(4)
    w, h, d = 21, 9, 6                                                 # W not a multiple of 4
    f = rng.uniform(-0.004, 0.016, (d + 2, h + 2, w + 2)).astype(np.float32)
    f[3, 2, 3] = np.nan; f[3, 2, 4] = 0.01; f[3, 2, 5] = 0.0; f[3, 2, 6] = np.inf; f[3, 2, 7] = -np.inf
    obs = (rng.uniform(size=f.shape) < 0.2).astype(np.float32)
    with Simulation(w, h, d, 0) as sim:
        sim.set("dens", f); sim.set("obs", obs)
        for z in range(d + 2):
            for alpha in (0.2, 0.0, 0.75):
                assert np.array_equal(sim.render_slice("dens", z, obstacle_alpha=alpha), opp.render_slice(f, obs, z, 0.0, 0.01, alpha))
        viewer.write_ppm(tmp_path / "s.ppm", sim.render_slice("dens", 3))
        raw = (tmp_path / "s.ppm").read_bytes()
        assert raw.startswith(b"P6\n23 11\n255\n") and len(raw) == 13 + 23 * 11 * 3
        with pytest.raises(Exception, match="bad argument"):
            sim.render_slice("dens", d + 2)


@pytest.mark.gpu
def test_driver_and_headless_viewer_outputs(tmp_path):
    """fluid_tunnel --view-* (C++ host) and `python -m fluid_simulation_b200.viewer` on the frames it wrote:
    both must give what the checker computes from the CPU solver's last frame."""
    import oracle
    from fluid_simulation_b200 import masks
    w, h, d, acc, iters = 40, 24, 24, 6, 5
    out = tmp_path / "data"; out.mkdir()
    tunnel = os.path.join(ROOT, "fluid_simulation_b200", "host", "fluid_tunnel")
    res = subprocess.run([tunnel, "--w", str(w), "--h", str(h), "--d", str(d), "--iter", str(iters), "--acc", str(acc),
                          "--obstacle", "sphere", "--out", str(out), "--view-ppm", str(tmp_path / "a.ppm"), "--view-field", "v_x",
                          "--view-z", "11", "--view-streamlines", str(tmp_path / "a.txt")], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    cpu = oracle.OracleSim(w, h, d, acc=acc, order=oracle.ORDER_RB)
    cpu.set("obs", masks.sphere(w, h, d)); cpu.run_iterations(iters)
    st = {k: cpu.get(k) for k in ("v_x", "v_y", "v_z", "obs", "dens")}
    head = b"P6\n%d %d\n255\n" % (w + 2, h + 2)
    assert (tmp_path / "a.ppm").read_bytes() == head + opp.render_slice(st["v_x"], st["obs"], 11, -10.0, 10.0).tobytes()
    lines, norm = opp.streamlines(st["v_x"], st["v_y"], st["v_z"], st["obs"])
    txt = (tmp_path / "a.txt").read_text().split("\n")
    got_len = [int(l.split()[1]) for l in txt if l.startswith("line")]
    got_norm = [float(l.split()[2]) for l in txt if l.startswith("line")]
    got_pts = np.array([[float(v) for v in l.split()] for l in txt if l and not l.startswith("line")]).reshape(-1, 3)
    assert got_len == [len(l) for l in lines] and np.array_equal(np.array(got_norm), norm)
    if lines:
        assert np.array_equal(got_pts, np.concatenate(lines) - 1)                     # %.17g round-trips float64
    # the Python headless viewer on the same frame files
    res = subprocess.run([sys.executable, "-m", "fluid_simulation_b200.viewer", str(out), "--field", "dens", "--z", "13",
                          "--ppm", str(tmp_path / "b.ppm"), "--streamlines", str(tmp_path / "b.npz")],
                         capture_output=True, text=True, cwd=ROOT)
    assert res.returncode == 0, res.stderr
    assert (tmp_path / "b.ppm").read_bytes() == head + opp.render_slice(st["dens"], st["obs"], 13, 0.0, 0.01).tobytes()
    z = np.load(tmp_path / "b.npz")
    assert list(z["line_len"]) == [len(l) for l in lines] and np.array_equal(z["norm_speed"], norm)
    assert z["rgba"].shape == (len(lines), 4)
    if lines:
        assert np.array_equal(z["points"], np.concatenate(lines) - 1)
