"""GPU parity tests proper: libfluid_cuda (through its C ABI, via ctypes) against the CPU oracle on
the same seeded inputs.  Bar: BIT-EXACT float32 for every function -- the kernels are compiled with
--fmad=false and keep the reference's operand order (north_star allows 1e-5 relative L2; we assert
bitwise equality and would only fall back to that tolerance if a platform libm differed).

Sweep orderings: FLUID_SOLVER_RB / JACOBI are compared with the ordering-matched CPU variants,
FLUID_SOLVER_LEX with the reference's own lexicographic order (oracle ORDER_LEX, itself pinned
bit-for-bit to the reference object code in test_oracle_vs_reference.py).
"""
import numpy as np
import pytest

from oracle import OracleSim, FIELD_NAMES, ORDER_LEX, ORDER_RB, ORDER_JACOBI
from tests.util import assert_bitwise, rel_l2

pytestmark = pytest.mark.gpu

SIZES = [(24, 16, 12), (17, 9, 5), (64, 32, 16), (5, 3, 2), (1, 1, 1), (33, 1, 7)]
SOLVERS = [("rb", 0x0, ORDER_RB), ("jacobi", 0x1, ORDER_JACOBI), ("lex", 0x2, ORDER_LEX)]


def make_pair(w, h, d, acc, flag, order, seed=3, obs=None, random=True, **kw):
    from fluid_simulation_b200 import Simulation, masks
    gpu = Simulation(w, h, d, 1, acc=acc, flags=flag, **kw)
    cpu = OracleSim(w, h, d, acc=acc, order=order, **kw)
    if random:
        st = masks.random_state(w, h, d, seed=seed, obs=obs)
    else:
        st = {"obs": masks.sphere(w, h, d) if obs is None else obs}
    gpu.load_state(st); cpu.load_state(st)
    return gpu, cpu


def check(gpu, cpu, tag, names=FIELD_NAMES):
    assert_bitwise(gpu.state(), cpu.state(), names, tag)


@pytest.mark.parametrize("w,h,d", SIZES)
@pytest.mark.parametrize("b", [0, 1, 2, 3])
def test_set_bounds(w, h, d, b):
    gpu, cpu = make_pair(w, h, d, 4, 0, ORDER_RB)
    gpu.setBounds(b, "v_y"); cpu.set_bounds(b, "v_y")
    check(gpu, cpu, f"setBounds b={b}")


@pytest.mark.parametrize("w,h,d", SIZES)
@pytest.mark.parametrize("b", [0, 1, 2, 3])
def test_advect(w, h, d, b):
    gpu, cpu = make_pair(w, h, d, 4, 0, ORDER_RB)
    f, p = [("dens", "buffer"), ("v_x", "v_x_prev"), ("v_y", "v_y_prev"), ("v_z", "v_z_prev")][b]
    # scale velocities so back-traces span many cells and hit the clamps
    for n in ("v_x", "v_y", "v_z", "v_x_prev", "v_y_prev", "v_z_prev"):
        a = cpu.get(n) * 30.0
        gpu.set(n, a); cpu.set(n, a)
    gpu.advect(b, f, p); cpu.advect(b, f, p)
    check(gpu, cpu, f"advect b={b}")


@pytest.mark.parametrize("w,h,d", SIZES + [(130, 70, 9), (256, 40, 6)])
@pytest.mark.parametrize("obstacle", [False, True])
def test_advect_velocity_fused(w, h, d, obstacle):
    """step()'s three velocity advections (simulation.cpp:125-127) as ONE launch against the three calls on the
    CPU: each reads what the previous one (and its setBounds) left at its own cell."""
    gpu, cpu = make_pair(w, h, d, 4, 0, ORDER_RB)
    if obstacle:
        from fluid_simulation_b200 import masks
        obs = masks.sphere(w, h, d)
        gpu.set("obs", obs); cpu.set("obs", obs)
    for n in ("v_x", "v_y", "v_z", "v_x_prev", "v_y_prev", "v_z_prev"):
        a = cpu.get(n) * 30.0
        gpu.set(n, a); cpu.set(n, a)
    gpu.advect_velocity()
    cpu.advect(1, "v_x", "v_x_prev"); cpu.advect(2, "v_y", "v_y_prev"); cpu.advect(3, "v_z", "v_z_prev")
    check(gpu, cpu, "fused velocity advection")


def test_staged_upload_equals_set():
    """stage_field + set_staged (H2D from a pinned copy of the held planes) leaves the same device state as set()."""
    from fluid_simulation_b200 import masks
    gpu, cpu = make_pair(40, 24, 12, 3, 0, ORDER_RB)
    obs = masks.sphere(40, 24, 12)
    v = cpu.get("v_y") * 2.0
    gpu.set_staged("obs", gpu.stage_field(obs)); cpu.set("obs", obs)
    gpu.set_staged("v_y", gpu.stage_field(v)); cpu.set("v_y", v)
    gpu.run_iterations(1); cpu.run_iterations(1)
    check(gpu, cpu, "state after staged uploads")


@pytest.mark.parametrize("w,h,d", SIZES)
def test_divergence_and_gradient(w, h, d):
    gpu, cpu = make_pair(w, h, d, 4, 0, ORDER_RB)
    gpu.divergence(); cpu.divergence()
    check(gpu, cpu, "divergence")
    st = cpu.state()
    st["pressure"] = np.random.default_rng(5).uniform(-1, 1, cpu.shape).astype(np.float32)
    gpu.set("pressure", st["pressure"]); cpu.set("pressure", st["pressure"])
    gpu.gradient(); cpu.gradient()
    check(gpu, cpu, "gradient")


@pytest.mark.parametrize("w,h,d", [(16, 16, 16), (64, 32, 16), (32, 8, 2), (16, 16, 8), (16, 8, 8), (64, 64, 128),
                                   (24, 16, 12), (100, 10, 10)])
def test_gradient_extreme_pressures(w, h, d):
    """Where h = 1/cbrtf(W*H*D) is a power of two (the first three shapes) k_gradient multiplies by 1/h
    instead of dividing by h; the claim is that this is exact for EVERY float, so the pressure field holds
    zeros of both signs, denormals, the smallest / largest normals and values whose differences overflow.
    The other shapes have a non-power-of-two h: the kernel multiplies by the rounded reciprocal and corrects with
    one FMA (csrc/recip_div.h) inside [2^-60, 2^60] and divides outside.  (16, 8, 8) and (64, 64, 128) have the
    significand of h that 256 x 256 x 512 and 1024 x 512 x 512 have (W*H*D = 2^(3k+1): h = 2^-k / cbrt(2)).)"""
    gpu, cpu = make_pair(w, h, d, 4, 0, ORDER_RB)
    rng = np.random.default_rng(w * 131 + d)
    shape = cpu.shape
    mant = rng.integers(0, 1 << 23, shape, dtype=np.uint32)
    sign = rng.integers(0, 2, shape, dtype=np.uint32) << 31
    kind = rng.integers(0, 8, shape)
    expo = np.select([kind == 0, kind == 1, kind == 2, kind == 3, kind == 4],
                     [0, 1, 254, rng.integers(1, 30, shape), rng.integers(225, 255, shape)],
                     rng.integers(100, 150, shape)).astype(np.uint32)
    bits = sign | (expo << 23) | np.where(kind == 5, 0, mant).astype(np.uint32)     # kind 0: denormals / zeros, 5: powers of two
    bits[kind == 6] = np.where(rng.integers(0, 2, int((kind == 6).sum())) == 0, 0x00000000, 0x80000000)   # +0 / -0
    p = bits.view(np.float32)
    assert np.isfinite(p).all() and (np.abs(p[p != 0]) < np.finfo(np.float32).tiny).any()
    gpu.set("pressure", p); cpu.set("pressure", p)
    for n in ("v_x", "v_y", "v_z"):            # zero velocity in half of the tunnel: there the result is -grad p itself
        v = cpu.get(n); v[: shape[0] // 2] = 0.0
        gpu.set(n, v); cpu.set(n, v)
    gpu.gradient(); cpu.gradient()
    got, want = gpu.state(), cpu.state()
    tiny = np.finfo(np.float32).tiny
    assert ((np.abs(want["v_x"]) < tiny) & (want["v_x"] != 0)).any()      # denormal quotients were exercised
    for n in ("v_x", "v_y", "v_z"):
        a, b = got[n], want[n]
        nan = np.isnan(b)                      # inf - inf: NaN on both sides, payloads are not compared
        assert np.array_equal(np.isnan(a), nan), n
        assert np.array_equal(a.view(np.uint32)[~nan], b.view(np.uint32)[~nan]), n
    assert np.isinf(want["v_x"]).any()         # overflowing differences were exercised


@pytest.mark.parametrize("name,flag,order", SOLVERS)
@pytest.mark.parametrize("w,h,d", SIZES)
@pytest.mark.parametrize("b", [0, 2])
def test_linear_solver(name, flag, order, w, h, d, b):
    gpu, cpu = make_pair(w, h, d, 5, flag, order)
    gpu.linearSolver(b, "v_y", "v_y_prev", 0.37, 3.22); cpu.linear_solver(b, "v_y", "v_y_prev", 0.37, 3.22)
    check(gpu, cpu, f"linearSolver[{name}] b={b}")


@pytest.mark.parametrize("name,flag,order", SOLVERS)
def test_diffuse_project(name, flag, order):
    gpu, cpu = make_pair(24, 16, 12, 6, flag, order)
    gpu.diffuse(1, "v_x", "v_x_prev"); cpu.diffuse(1, "v_x", "v_x_prev")
    check(gpu, cpu, f"diffuse[{name}]")
    gpu.project(); cpu.project()
    check(gpu, cpu, f"project[{name}]")


@pytest.mark.parametrize("name,flag,order", SOLVERS)
@pytest.mark.parametrize("w,h,d,acc,steps", [(24, 16, 12, 7, 5), (64, 64, 64, 20, 3), (40, 20, 10, 4, 8)])
def test_run_iterations_from_rest(name, flag, order, w, h, d, acc, steps):
    """The reference's own start: zero fields, inlet only (simulation.cpp:38-43, 63-71)."""
    gpu, cpu = make_pair(w, h, d, acc, flag, order, random=False)
    gpu.run_iterations(steps); cpu.run_iterations(steps)
    check(gpu, cpu, f"{steps} run() iterations [{name}]")
    s = gpu.stats()
    assert np.isclose(s["dens_sum"], cpu.get("dens").astype(np.float64).sum(), rtol=1e-12)
    assert s["vx_max"] == cpu.get("v_x").max() and s["vz_min"] == cpu.get("v_z").min()


@pytest.mark.parametrize("name,flag,order", SOLVERS)
def test_run_iterations_random_state(name, flag, order):
    gpu, cpu = make_pair(20, 12, 9, 5, flag, order)
    gpu.run_iterations(3); cpu.run_iterations(3)
    check(gpu, cpu, f"3 iterations from a random state [{name}]")


def test_elide_dead_is_bit_identical():
    """Skipping diffuse(0,dens,buffer) (simulation.cpp:135) cannot change any output."""
    from fluid_simulation_b200 import Simulation, masks, ELIDE_DEAD
    w, h, d = 24, 16, 12
    a = Simulation(w, h, d, 1, acc=6); b = Simulation(w, h, d, 1, acc=6, flags=ELIDE_DEAD)
    obs = masks.sphere(w, h, d)
    a.set("obs", obs); b.set("obs", obs)
    a.run_iterations(4); b.run_iterations(4)
    assert_bitwise(a.state(), b.state(), [n for n in FIELD_NAMES], "ELIDE_DEAD")


def test_sparse_mutators_match_reference_api():
    gpu, cpu = make_pair(12, 10, 8, 3, 0, ORDER_RB)
    gpu.addObstacle(3, 4, 5); cpu.add_obstacle(3, 4, 5)
    gpu.addDensity(2, 2, 2, 0.25); cpu.add_density(2, 2, 2, 0.25)
    gpu.setVelocity(4, 5, 6, 1.0, -2.0, 3.0); cpu.set_velocity(4, 5, 6, 1.0, -2.0, 3.0)
    gpu.run_iterations(1); cpu.run_iterations(1)
    check(gpu, cpu, "mutators + 1 iteration")


def test_invariants_large_grid():
    """Size-independent properties at a BASELINE-sized grid (256x128x128, acc=8, airfoil):
    solids stay 0, edge/corner ghosts stay 0, x=W+1 ghost equals x=W, scalars mirror at walls."""
    from fluid_simulation_b200 import Simulation, masks
    w, h, d = 256, 128, 128
    sim = Simulation(w, h, d, 1, acc=8)
    obs = masks.airfoil(w, h, d)
    sim.set("obs", obs)
    sim.run_iterations(2)
    solid = obs == 1.0
    for n in ("v_x", "v_y", "v_z", "dens"):
        a = sim.get(n)
        assert np.all(a[solid] == 0.0), n
        assert np.all(np.isfinite(a)), n
        assert np.all(a[1:-1, 1:-1, -1] == a[1:-1, 1:-1, -2]), f"{n}: outflow ghost"
        for e in (a[0, 0, :], a[0, -1, :], a[-1, 0, :], a[-1, -1, :], a[0, :, 0], a[0, :, -1],
                  a[-1, :, 0], a[-1, :, -1], a[:, 0, 0], a[:, 0, -1], a[:, -1, 0], a[:, -1, -1]):
            assert np.all(e == 0.0), f"{n}: edge/corner ghost written"
    vy = sim.get("v_y")
    assert np.all(vy[1:-1, 0, 1:-1] == -vy[1:-1, 1, 1:-1])
    dens = sim.get("dens")
    assert np.all(dens[1:-1, 0, 1:-1] == dens[1:-1, 1, 1:-1])
    assert dens.max() > 0


def test_batched_mutators_with_repeated_cells():
    """A batch that names a cell more than once behaves like the reference's sequence of calls
    (simulation.cpp:155-178): densities add up in call order (float rounding included), the last
    setVelocity wins."""
    gpu, cpu = make_pair(12, 10, 8, 3, 0, ORDER_RB)
    rng = np.random.default_rng(9)
    cells = np.array([[2, 3, 4], [5, 5, 5], [2, 3, 4], [7, 1, 2], [2, 3, 4], [5, 5, 5]], dtype=np.int32)
    amounts = np.array([1e-8, 0.3, 0.7, 0.25, 1e8, -0.3], dtype=np.float32)          # order-sensitive in float32
    vel = rng.uniform(-1, 1, (len(cells), 3)).astype(np.float32)
    gpu.add_density_cells(np.ascontiguousarray(cells), amounts)
    gpu.set_velocity_cells(np.ascontiguousarray(cells), np.ascontiguousarray(vel))
    for (x, y, z), a in zip(cells, amounts):
        cpu.add_density(int(x), int(y), int(z), float(a))
    for (x, y, z), v in zip(cells, vel):
        cpu.set_velocity(int(x), int(y), int(z), float(v[0]), float(v[1]), float(v[2]))
    check(gpu, cpu, "repeated cells in one batch")
