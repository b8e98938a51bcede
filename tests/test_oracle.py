"""CPU tests of the checker itself: the plain-C restatement against (a) the committed golden
vectors produced by the reference object code and (b) the reference itself when oracle/_ref is
present.  Bit-exact."""
import numpy as np
import pytest

from oracle import OracleSim, RefSim, have_ref, FIELD_NAMES, ORDER_LEX, ORDER_RB, ORDER_JACOBI
from fluid_simulation_b200 import masks
from tests.golden_util import golden_cases, load_golden, replay
from tests.util import assert_bitwise, rel_l2


def oracle_api(o):
    return dict(run=o.run_iterations, bounds=o.set_bounds, linsolve=o.linear_solver, diffuse=o.diffuse,
                project=o.project, advect=o.advect, step=o.step)


@pytest.mark.parametrize("name", golden_cases())
def test_oracle_matches_golden(name):
    meta, inputs, outputs = load_golden(name)
    o = OracleSim(meta["w"], meta["h"], meta["d"], acc=meta["acc"], order=ORDER_LEX)
    o.load_state(inputs)
    replay(o, meta, oracle_api(o))
    assert_bitwise(o.state(), outputs, list(outputs), f"oracle vs golden {name}")


needs_ref = pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference)")


@needs_ref
@pytest.mark.parametrize("w,h,d", [(24, 16, 12), (7, 5, 3), (1, 1, 1), (16, 1, 9)])
def test_oracle_matches_reference_per_function(w, h, d):
    st = masks.random_state(w, h, d, seed=w)
    r = RefSim(w, h, d, acc=5, threads=1); o = OracleSim(w, h, d, acc=5)
    r.load_state(st); o.load_state(st)
    for b in range(4):
        r.set_bounds(b, "v_x"); o.set_bounds(b, "v_x")
    assert_bitwise(o.state(), r.state(), tag="setBounds")
    r.linear_solver(3, "v_z", "v_z_prev", 0.9, 6.4); o.linear_solver(3, "v_z", "v_z_prev", 0.9, 6.4)
    assert_bitwise(o.state(), r.state(), tag="linearSolver")
    r.diffuse(0, "dens", "buffer"); o.diffuse(0, "dens", "buffer")
    assert_bitwise(o.state(), r.state(), tag="diffuse")
    r.project(); o.project()
    assert_bitwise(o.state(), r.state(), tag="project")
    for b, (f, p) in enumerate((("dens", "buffer"), ("v_x", "v_x_prev"), ("v_y", "v_y_prev"), ("v_z", "v_z_prev"))):
        r.advect(b, f, p); o.advect(b, f, p)
    assert_bitwise(o.state(), r.state(), tag="advect")


@needs_ref
def test_oracle_matches_reference_run_1_thread():
    """N run() iterations from rest with a sphere, reference at OMP_NUM_THREADS=1."""
    w, h, d = 32, 24, 16
    obs = masks.sphere(w, h, d)
    r = RefSim(w, h, d, acc=8, threads=1); o = OracleSim(w, h, d, acc=8)
    r.set("obs", obs); o.set("obs", obs)
    r.run_iterations(6); o.run_iterations(6)
    assert_bitwise(o.state(), r.state(), tag="6 iterations")


def test_elide_dead_density_diffusion_is_dead():
    """diffuse(0,dens,buffer) is overwritten by advect -- the oracle proves it bitwise."""
    w, h, d = 20, 12, 10
    obs = masks.sphere(w, h, d)
    a = OracleSim(w, h, d, acc=6); b = OracleSim(w, h, d, acc=6, elide_dead=True)
    a.set("obs", obs); b.set("obs", obs)
    a.run_iterations(4); b.run_iterations(4)
    assert_bitwise(a.state(), b.state(), tag="elide dead")


@pytest.mark.parametrize("order", [ORDER_RB, ORDER_JACOBI])
def test_ordering_variants_converge_to_the_same_solution(order):
    """The ordering-matched variants solve the same linear system: with many sweeps on a
    diagonally dominant diffusion problem they agree with lexicographic GS to ~1e-6."""
    w, h, d = 12, 10, 8
    st = masks.random_state(w, h, d, seed=2, obs=masks.empty(w, h, d))
    lex = OracleSim(w, h, d, acc=200); var = OracleSim(w, h, d, acc=200, order=order)
    lex.load_state(st); var.load_state(st)
    lex.linear_solver(0, "dens", "buffer", 0.3, 2.8); var.linear_solver(0, "dens", "buffer", 0.3, 2.8)
    assert rel_l2(var.get("dens"), lex.get("dens")) < 1e-5


def test_set_bounds_keeps_edges_and_prezero_ghosts():
    w, h, d = 6, 5, 4
    o = OracleSim(w, h, d, acc=1)
    f = np.arange(o.shape[0] * o.shape[1] * o.shape[2], dtype=np.float32).reshape(o.shape) + 1
    obs = masks.empty(w, h, d); obs[2, 1, 3] = 1.0            # solid touching the y=1 wall
    o.set("obs", obs); o.set("v_y", f)
    o.set_bounds(2, "v_y")
    g = o.get("v_y")
    assert g[2, 1, 3] == 0.0 and g[2, 0, 3] == -f[2, 1, 3]    # ghost keeps the PRE-zero value
    assert g[0, 0, 0] == f[0, 0, 0] and g[0, 0, 3] == f[0, 0, 3]   # corner and edge untouched
    assert g[2, 1, 2] == 0.0                                   # no-slip neighbour zeroed
