"""Stated-tolerance tier: the default red-black (and the Jacobi) trajectories against the reference's own
lexicographic ordering after N steps -- velocity, density and the post-projection divergence residual.
Tolerances, their calibration against the reference's 1-thread-vs-8-thread spread and the metric definitions
are in tests/tolerance.py.  The GPU tests go through the C ABI (fluid_step, fluid_call(DIVERGENCE),
fluid_divergence_norm); the CPU test checks the same bounds on the ordering-matched restatements, which the
GPU is bit-identical to, so the stated numbers are validated even where no GPU is at hand."""
import numpy as np
import pytest

from oracle import OracleSim, ORDER_RB, ORDER_JACOBI
from tests import tolerance as T

ORDERS = [("rb", 0x0, ORDER_RB), ("jacobi", 0x1, ORDER_JACOBI)]


def _masks():
    from fluid_simulation_b200 import masks
    return masks


# (tag, w, h, d, acc, mask name, steps): configs[0] = 64^3 empty tunnel acc 20; configs[1] = 128^3-plane sphere
# acc 40 at reduced depth
GPU_CASES = [("configs[0] 64^3 empty", 64, 64, 64, 20, "empty", 5),
             ("configs[0] 64^3 empty", 64, 64, 64, 20, "empty", 20),
             ("configs[1] 128x128x24 sphere", 128, 128, 24, 40, "sphere", 5)]


@pytest.mark.gpu
@pytest.mark.parametrize("name,flag,order", ORDERS)
@pytest.mark.parametrize("tag,w,h,d,acc,mask,steps", GPU_CASES)
def test_gpu_trajectory_within_stated_tolerance(name, flag, order, tag, w, h, d, acc, mask, steps):
    from fluid_simulation_b200 import Simulation
    obs = getattr(_masks(), mask)(w, h, d)
    ref, how = T.reference_sim(w, h, d, acc)
    gpu = Simulation(w, h, d, steps, acc=acc, flags=flag)
    ref.set("obs", obs); gpu.set("obs", obs)
    ref.run_iterations(steps); gpu.run_iterations(steps)
    rs, gs = ref.state(), gpu.state()
    div_ref = T.divergence_residual_cpu(rs, w, h, d, acc)
    gpu.divergence()                                   # div of the final velocity, on the device (:295-319)
    div_gpu, div_max = gpu.divergence_norm()           # fluid_divergence_norm: L2 and max over fluid cells
    assert div_gpu == pytest.approx(T.divergence_residual_cpu(gs, w, h, d, acc), rel=1e-6)
    dist, ratio = T.check(name, gs, rs, div_gpu, div_ref, T.TOL, f"{tag}, {steps} steps vs {how}")
    print(f"{tag} {steps} steps [{name}] vs {how}: {dist}, divergence residual ratio {ratio:.3f}")


@pytest.mark.gpu
@pytest.mark.parametrize("name,flag,order", ORDERS)
@pytest.mark.parametrize("w,h,d,acc,mask", [(64, 64, 64, 20, "empty"), (128, 128, 24, 40, "sphere")])
def test_gpu_one_step_from_identical_state(name, flag, order, w, h, d, acc, mask):
    """Both sides start from the SAME developed state (4 reference steps), then take one step."""
    from fluid_simulation_b200 import Simulation
    obs = getattr(_masks(), mask)(w, h, d)
    ref, how = T.reference_sim(w, h, d, acc)
    ref.set("obs", obs)
    ref.run_iterations(4)
    gpu = Simulation(w, h, d, 1, acc=acc, flags=flag)
    gpu.load_state(ref.state())
    ref.run_iterations(1); gpu.run_iterations(1)
    rs, gs = ref.state(), gpu.state()
    gpu.divergence()
    div_gpu, _ = gpu.divergence_norm()
    T.check(name, gs, rs, div_gpu, T.divergence_residual_cpu(rs, w, h, d, acc), T.TOL_ONE_STEP,
            f"{w}x{h}x{d} one step from an identical state vs {how}")


@pytest.mark.parametrize("name,flag,order", ORDERS)
def test_ordering_matched_restatement_within_stated_tolerance(name, flag, order):
    """CPU: the ordering-matched loops (what the GPU reproduces bit for bit) obey the same stated bounds
    against the reference's ordering -- 64^3 empty tunnel, acc 20, 5 steps (configs[0], shortened)."""
    w, h, d, acc, steps = 64, 64, 64, 20, 5
    obs = _masks().empty(w, h, d)
    ref, how = T.reference_sim(w, h, d, acc)
    sim = OracleSim(w, h, d, acc=acc, order=order)
    ref.set("obs", obs); sim.set("obs", obs)
    ref.run_iterations(steps); sim.run_iterations(steps)
    rs, ss = ref.state(), sim.state()
    T.check(name, ss, rs, T.divergence_residual_cpu(ss, w, h, d, acc), T.divergence_residual_cpu(rs, w, h, d, acc),
            T.TOL, f"restatement, {steps} steps vs {how}")
    # and it is NOT trivially loose: the orderings really differ
    assert T.distances(ss, rs)["v_x"] > 1e-3
