"""fluid_simulation_b200 -- B200 (sm_100a) drop-in for the reference wind-tunnel solver's hot path.

`Simulation` mirrors the reference class (simulation.h) and forwards to libfluid_cuda.so
(include/fluid_cuda.h).  Importing the package does not need a GPU; constructing a `Simulation`
does, and fails loudly when the CUDA library or device is missing (there is no CPU path).
"""
from .simulation import (Simulation, FluidError, FIELDS, FIELD_NAMES, SOLVER_RB, SOLVER_JACOBI,
                         SOLVER_LEX, ELIDE_DEAD, NO_TILED, TILED_ALWAYS, VOX_HASHED, VOX_FIXED, VOX_CELLS)
from . import masks, stl

__all__ = ["Simulation", "FluidError", "FIELDS", "FIELD_NAMES", "SOLVER_RB", "SOLVER_JACOBI",
           "SOLVER_LEX", "ELIDE_DEAD", "NO_TILED", "TILED_ALWAYS", "VOX_HASHED", "VOX_FIXED", "VOX_CELLS",
           "masks", "stl"]
