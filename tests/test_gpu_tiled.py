"""The TMA-fed register-streaming sweep kernel (csrc/sweep_stream.cuh) against the ordering-matched
CPU oracle, bit-exact, over warp-column, row-tile, ragged and multi-chunk shapes; and against the
simple global-memory kernels (FLUID_NO_TILED) on a larger grid.  Shapes with W % 4 != 0 or W > 1024
exercise the documented hand-over to the simple kernels under FLUID_TILED_ALWAYS."""
import os

import numpy as np
import pytest

from oracle import OracleSim, FIELD_NAMES, ORDER_RB, ORDER_JACOBI
from tests.util import assert_bitwise

pytestmark = pytest.mark.gpu

TILED_ALWAYS, NO_TILED = 0x40, 0x20
SOLVERS = [("rb", 0x0, ORDER_RB), ("jacobi", 0x1, ORDER_JACOBI)]
# W around the 128-cell warp columns (1, 2, 4, 8 columns), H around the row-tile heights (30 / 14 /
# 6 / 2 owned rows), D around chunk sizes
SHAPES = [(130, 33, 9), (128, 30, 8), (127, 31, 17), (257, 61, 12), (100, 29, 5), (24, 16, 12),
          (5, 3, 2), (1, 1, 1), (260, 8, 4), (129, 90, 3), (256, 40, 20), (512, 13, 9), (384, 20, 7),
          (64, 64, 33), (132, 70, 6), (8, 2, 2), (4, 1, 1), (252, 17, 5), (516, 6, 3),
          (1024, 9, 4), (772, 5, 6), (1028, 4, 3)]


def pair(w, h, d, acc, flag, order, seed=7):
    from fluid_simulation_b200 import Simulation, masks
    gpu = Simulation(w, h, d, 1, acc=acc, flags=flag | TILED_ALWAYS)
    cpu = OracleSim(w, h, d, acc=acc, order=order)
    st = masks.random_state(w, h, d, seed=seed)
    gpu.load_state(st); cpu.load_state(st)
    return gpu, cpu


@pytest.mark.parametrize("name,flag,order", SOLVERS)
@pytest.mark.parametrize("w,h,d", SHAPES)
@pytest.mark.parametrize("b,acc", [(0, 3), (1, 2), (2, 5), (3, 4)])
def test_tiled_linear_solver(name, flag, order, w, h, d, b, acc):
    gpu, cpu = pair(w, h, d, acc, flag, order)
    gpu.linearSolver(b, "v_y", "v_y_prev", 0.37, 3.22); cpu.linear_solver(b, "v_y", "v_y_prev", 0.37, 3.22)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, f"tiled linearSolver[{name}] {w}x{h}x{d} b={b}")


@pytest.mark.parametrize("zc", [1, 2, 3, 7])
@pytest.mark.parametrize("name,flag,order", SOLVERS)
def test_tiled_chunk_sizes(monkeypatch, zc, name, flag, order):
    monkeypatch.setenv("FLUID_TILED_ZC", str(zc))
    gpu, cpu = pair(131, 35, 11, 4, flag, order)
    gpu.linearSolver(3, "v_z", "v_z_prev", 1.0, 6.0); cpu.linear_solver(3, "v_z", "v_z_prev", 1.0, 6.0)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, f"tiled[{name}] zc={zc}")


@pytest.mark.parametrize("name,flag,order", SOLVERS)
def test_tiled_full_steps(name, flag, order):
    from fluid_simulation_b200 import Simulation, masks
    w, h, d, acc = 136, 40, 24, 6
    gpu = Simulation(w, h, d, 1, acc=acc, flags=flag)          # W >= 96: tiled path by default
    cpu = OracleSim(w, h, d, acc=acc, order=order)
    obs = masks.sphere(w, h, d)
    gpu.set("obs", obs); cpu.set("obs", obs)
    gpu.run_iterations(3); cpu.run_iterations(3)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, f"tiled 3 iterations [{name}]")


@pytest.mark.parametrize("name,flag,order", SOLVERS)
@pytest.mark.parametrize("w,h,d", [(320, 24, 16), (512, 20, 12), (64, 40, 20)])
def test_tiled_full_steps_other_column_counts(name, flag, order, w, h, d):
    """Whole run() iterations through the four-warp-column (256 < W <= 512) and one-warp-column (W <= 128)
    kernels: the 4-field diffusion launch, the v <-> v_prev swap and the dependent launches on those shapes."""
    from fluid_simulation_b200 import Simulation, masks
    acc = 5
    gpu = Simulation(w, h, d, 1, acc=acc, flags=flag)
    cpu = OracleSim(w, h, d, acc=acc, order=order)
    obs = masks.box(w, h, d)
    gpu.set("obs", obs); cpu.set("obs", obs)
    gpu.run_iterations(2); cpu.run_iterations(2)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, f"tiled 2 iterations {w}x{h}x{d} [{name}]")


@pytest.mark.parametrize("name,flag,order", SOLVERS)
def test_tiled_equals_simple_kernels_256(name, flag, order):
    """A/B at a BASELINE-sized plane (256x256x48, airfoil): tiled vs global-memory kernels."""
    from fluid_simulation_b200 import Simulation, masks
    w, h, d, acc = 256, 256, 48, 10
    obs = masks.airfoil(w, h, d)
    a = Simulation(w, h, d, 1, acc=acc, flags=flag)
    b = Simulation(w, h, d, 1, acc=acc, flags=flag | NO_TILED)
    a.set("obs", obs); b.set("obs", obs)
    a.run_iterations(2); b.run_iterations(2)
    assert_bitwise(a.state(), b.state(), FIELD_NAMES, f"tiled vs simple [{name}]")


@pytest.mark.parametrize("name,flag,order", SOLVERS)
def test_tiled_equals_simple_kernels_wide(name, flag, order):
    """A/B at the W = 1024 plane of BASELINE config 5 (eight warp columns, two owned rows)."""
    from fluid_simulation_b200 import Simulation, masks
    w, h, d, acc = 1024, 96, 20, 6
    obs = masks.sphere(w, h, d)
    a = Simulation(w, h, d, 1, acc=acc, flags=flag)
    b = Simulation(w, h, d, 1, acc=acc, flags=flag | NO_TILED)
    a.set("obs", obs); b.set("obs", obs)
    a.run_iterations(2); b.run_iterations(2)
    assert_bitwise(a.state(), b.state(), FIELD_NAMES, f"tiled vs simple, W=1024 [{name}]")


@pytest.mark.parametrize("name,flag,order", SOLVERS)
@pytest.mark.parametrize("d", [16, 40])
def test_headline_plane_vs_oracle(name, flag, order, d):
    """The 256 x 256 plane of BASELINE.json configs[2] against the CPU ORACLE (not against the repo's own
    simple kernels): airfoil mask, acc = 10, two run() iterations, default segment plan (no FLUID_TILED_ZC),
    so the balanced plan, the 4-field diffusion launch, the v <-> v_prev swap and the dependent launches of
    the benchmarked kernel configuration meet the oracle bit for bit."""
    from fluid_simulation_b200 import Simulation, masks
    w, h, acc = 256, 256, 10
    gpu = Simulation(w, h, d, 1, acc=acc, flags=flag)
    cpu = OracleSim(w, h, d, acc=acc, order=order)
    obs = masks.airfoil(w, h, d)
    gpu.set("obs", obs); cpu.set("obs", obs)
    gpu.run_iterations(2); cpu.run_iterations(2)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, f"256x256x{d} airfoil, 2 iterations [{name}]")


def test_sweep_trace_writes_per_cta_timelines(monkeypatch, tmp_path):
    """FLUID_SW_TRACE / FLUID_SW_TRACE_AT: the diagnostic that profiles/r02_slab_trace.txt was made with -- per CTA
    { start, first segment done, end } of the chosen sweep launches; the results are not touched."""
    from fluid_simulation_b200 import Simulation, masks
    monkeypatch.setenv("FLUID_SW_TRACE", str(tmp_path / "trace"))
    monkeypatch.setenv("FLUID_SW_TRACE_AT", "3,9")
    w, h, d, acc = 128, 40, 24, 6
    gpu = Simulation(w, h, d, 1, acc=acc, flags=0)
    cpu = OracleSim(w, h, d, acc=acc, order=ORDER_RB)
    obs = masks.sphere(w, h, d)
    gpu.set("obs", obs); cpu.set("obs", obs)
    gpu.run_iterations(1); cpu.run_iterations(1)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, "traced run")
    for launch in (3, 9):
        lines = (tmp_path / f"trace.r0.l{launch}.txt").read_text().splitlines()
        assert lines[0].startswith("# rank 0 launch")
        rows = [[int(v) for v in ln.split()] for ln in lines[1:]]
        assert rows and all(len(r) == 6 for r in rows)
        busy = [r for r in rows if r[3] < 10 ** 9]
        assert busy and all(r[1] <= r[3] for r in busy)              # start <= end, in ns from the first start


def test_headline_grid_vs_oracle_full_depth():
    """configs[2] at its FULL size (256^3, airfoil, red-black): one run() iteration at acc = 4 (the oracle's
    single thread needs ~40 s for it).  This is the regime the benchmark runs in and no smaller case reaches:
    148 CTAs, segments crossing column boundaries, z-ranges of ~130 planes in the 4-field launch."""
    from fluid_simulation_b200 import Simulation, masks
    w = h = d = 256
    acc = 4
    gpu = Simulation(w, h, d, 1, acc=acc, flags=0)
    cpu = OracleSim(w, h, d, acc=acc, order=ORDER_RB)
    obs = masks.airfoil(w, h, d)
    gpu.set("obs", obs); cpu.set("obs", obs)
    gpu.run_iterations(1); cpu.run_iterations(1)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, "256^3 airfoil, 1 iteration, acc 4 [rb]")


# shapes the temporally blocked kernel takes when FLUID_PAIR=1 (two sweeps per pass, csrc/sweep_pair*.cuh:
# 128 < W <= 256, W % 8 == 0, H even); odd sweep counts finish with one ordinary sweep
PAIR_SHAPES = [(136, 12, 7), (256, 24, 9), (200, 10, 5), (144, 2, 3), (256, 64, 33), (248, 30, 40), (160, 16, 1),
               (256, 256, 12)]


@pytest.mark.parametrize("unroll", ["2", "6"])
@pytest.mark.parametrize("w,h,d", PAIR_SHAPES)
@pytest.mark.parametrize("b,acc", [(0, 3), (1, 2), (2, 5), (3, 4)])
def test_pair_linear_solver(monkeypatch, w, h, d, b, acc, unroll):
    monkeypatch.setenv("FLUID_PAIR", "1"); monkeypatch.setenv("FLUID_PAIR_UNROLL", unroll)
    gpu, cpu = pair(w, h, d, acc, 0x0, ORDER_RB)
    gpu.linearSolver(b, "v_y", "v_y_prev", 0.37, 3.22); cpu.linear_solver(b, "v_y", "v_y_prev", 0.37, 3.22)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, f"pair-sweep linearSolver {w}x{h}x{d} b={b} acc={acc}")


@pytest.mark.parametrize("zc", [1, 2, 3, 7])
def test_pair_chunk_sizes(monkeypatch, zc):
    monkeypatch.setenv("FLUID_TILED_ZC", str(zc)); monkeypatch.setenv("FLUID_PAIR", "1")
    gpu, cpu = pair(256, 36, 11, 4, 0x0, ORDER_RB)
    gpu.linearSolver(3, "v_z", "v_z_prev", 1.0, 6.0); cpu.linear_solver(3, "v_z", "v_z_prev", 1.0, 6.0)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, f"pair-sweep zc={zc}")


def test_pair_equals_single_sweeps(monkeypatch):
    """A/B inside the library: FLUID_PAIR=1 routes the solves through k_sweep_pair (two sweeps per pass, with
    the 4-field diffusion launch, the buffer swap and an odd sweep count), the default through k_sweep_stream
    (one sweep per pass); whole run() iterations must come out bit-identical either way."""
    from fluid_simulation_b200 import Simulation, masks
    w, h, d, acc = 256, 128, 48, 7
    obs = masks.airfoil(w, h, d)
    monkeypatch.setenv("FLUID_PAIR", "1")
    a = Simulation(w, h, d, 1, acc=acc)
    a.set("obs", obs); a.run_iterations(2)
    monkeypatch.delenv("FLUID_PAIR")
    b = Simulation(w, h, d, 1, acc=acc)
    b.set("obs", obs); b.run_iterations(2)
    assert_bitwise(a.state(), b.state(), FIELD_NAMES, "pair vs single sweeps")


@pytest.mark.parametrize("name,flag,order", SOLVERS)
@pytest.mark.parametrize("w,h,d", [(512, 13, 9), (384, 20, 7), (260, 8, 4), (1024, 9, 4), (2052, 3, 5)])
def test_x_tiles(monkeypatch, name, flag, order, w, h, d):
    """Grids wider than a CTA are cut into x tiles with one redundantly computed halo quad beside every interior
    tile edge (default above W = 512; FLUID_SW_CFG=5 forces the tiled two-column CTA from W = 260 on)."""
    monkeypatch.setenv("FLUID_SW_CFG", "5")
    gpu, cpu = pair(w, h, d, 3, flag, order)
    gpu.linearSolver(1, "v_x", "v_x_prev", 0.37, 3.22); cpu.linear_solver(1, "v_x", "v_x_prev", 0.37, 3.22)
    assert_bitwise(gpu.state(), cpu.state(), FIELD_NAMES, f"x tiles [{name}] {w}x{h}x{d}")
