"""GPU vs the committed golden vectors of the reference (tests/golden, FLUID_SOLVER_LEX):
the CUDA path reproduces the reference object code's output bit for bit."""
import pytest

from tests.golden_util import golden_cases, load_golden, replay
from tests.util import assert_bitwise

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", golden_cases())
def test_cuda_lex_matches_reference_golden(name):
    from fluid_simulation_b200 import Simulation, SOLVER_LEX
    meta, inputs, outputs = load_golden(name)
    sim = Simulation(meta["w"], meta["h"], meta["d"], 1, acc=meta["acc"], flags=SOLVER_LEX)
    sim.load_state(inputs)
    api = dict(run=sim.run_iterations, bounds=sim.setBounds, linsolve=sim.linearSolver, diffuse=sim.diffuse,
               project=sim.project, advect=sim.advect, step=sim.step)
    replay(sim, meta, api)
    got = {k: sim.get(k) for k in outputs}
    assert_bitwise(got, outputs, list(outputs), f"CUDA LEX vs reference golden {name}")
