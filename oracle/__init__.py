"""oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes front-ends for the two CPU checkers:

* ``RefSim``    -- the UNMODIFIED reference (`/root/reference/simulation.cpp`) compiled in
                   place into ``oracle/_ref/libfluid_ref.so`` by ``oracle/Makefile``.
* ``OracleSim`` -- the plain-C restatement ``oracle/fluid_oracle.c`` (``liboracle.so``) with
                   the reference's lexicographic order plus the ordering-matched red-black
                   and Jacobi variants.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  ``fluid_simulation_b200`` never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

FIELD_IDS = {
    "v_x": 0, "v_y": 1, "v_z": 2, "v_x_prev": 3, "v_y_prev": 4, "v_z_prev": 5,
    "dens": 6, "buffer": 7, "obs": 8, "pressure": 9, "divergence": 10,
}
FIELD_NAMES = list(FIELD_IDS)
ORDER_LEX, ORDER_RB, ORDER_JACOBI = 0, 1, 2

_f, _i, _p = C.c_float, C.c_int, C.c_void_p


def ref_lib_path() -> str:
    return os.path.join(_HERE, "_ref", "libfluid_ref.so")


def oracle_lib_path() -> str:
    return os.path.join(_HERE, "liboracle.so")


def have_ref() -> bool:
    return os.path.exists(ref_lib_path())


def _fid(f) -> int:
    return FIELD_IDS[f] if isinstance(f, str) else int(f)


class _Base:
    """Shared numpy plumbing: fields are views of the C arrays, shape (D+2, H+2, W+2)."""

    def __init__(self, w, h, d):
        self.width, self.height, self.depth = w, h, d
        self.shape = (d + 2, h + 2, w + 2)

    def _view(self, ptr):
        n = self.shape[0] * self.shape[1] * self.shape[2]
        return np.ctypeslib.as_array(ptr, shape=(n,)).reshape(self.shape)

    def get(self, f) -> np.ndarray:
        return self.field(f).copy()

    def set(self, f, arr):
        self.field(f)[...] = np.asarray(arr, dtype=np.float32).reshape(self.shape)

    def state(self) -> dict:
        return {n: self.get(n) for n in FIELD_NAMES}

    def load_state(self, st: dict):
        for n, a in st.items():
            self.set(n, a)


class RefSim(_Base):
    """The reference's own `Simulation` (object code of /root/reference), via ref_harness.cpp."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            L = C.CDLL(ref_lib_path())
            L.ref_create.restype = _p
            L.ref_create.argtypes = [_i, _i, _i, _i, _i, _f, _f, _f, _i]
            L.ref_destroy.argtypes = [_p]
            L.ref_size.argtypes = [_p]; L.ref_size.restype = _i
            L.ref_field.argtypes = [_p, _i]; L.ref_field.restype = C.POINTER(_f)
            L.ref_set_acc.argtypes = [_p, _i]
            L.ref_set_threads.argtypes = [_i]
            L.ref_max_threads.restype = _i
            L.ref_add_obstacle.argtypes = [_p, _i, _i, _i]
            L.ref_add_density.argtypes = [_p, _i, _i, _i, _f]
            L.ref_set_velocity.argtypes = [_p, _i, _i, _i, _f, _f, _f]
            L.ref_set_bounds.argtypes = [_p, _i, _i]
            L.ref_linear_solver.argtypes = [_p, _i, _i, _i, _f, _f]
            L.ref_diffuse.argtypes = [_p, _i, _i, _i]
            L.ref_project.argtypes = [_p]
            L.ref_advect.argtypes = [_p, _i, _i, _i]
            L.ref_step.argtypes = [_p]
            L.ref_run_iteration.argtypes = [_p]
            L.ref_run_iterations.argtypes = [_p, _i]
            L.ref_load_stl.argtypes = [_p, C.c_char_p, _f, _f, _f, _f, _f, _f, _f]
            cls._lib = L
        return cls._lib

    def __init__(self, w, h, d, iter=1, speed=30, dt=0.05, diff=2.0e-5, visc=1.5e-5, acc=15, threads=1):
        super().__init__(w, h, d)
        L = self.lib()
        L.ref_set_threads(threads)
        self.h = L.ref_create(w, h, d, iter, speed, dt, diff, visc, acc)
        self.acc = acc

    def __del__(self):
        if getattr(self, "h", None):
            self.lib().ref_destroy(self.h)
            self.h = None

    @staticmethod
    def set_threads(n):
        RefSim.lib().ref_set_threads(n)

    @staticmethod
    def max_threads():
        return RefSim.lib().ref_max_threads()

    def field(self, f):
        return self._view(self.lib().ref_field(self.h, _fid(f)))

    def set_acc(self, acc):
        self.acc = acc
        self.lib().ref_set_acc(self.h, acc)

    def add_obstacle(self, x, y, z): self.lib().ref_add_obstacle(self.h, x, y, z)
    def add_density(self, x, y, z, a): self.lib().ref_add_density(self.h, x, y, z, a)
    def set_velocity(self, x, y, z, ax, ay, az): self.lib().ref_set_velocity(self.h, x, y, z, ax, ay, az)
    def set_bounds(self, b, f): self.lib().ref_set_bounds(self.h, b, _fid(f))
    def linear_solver(self, b, f, prev, a, c): self.lib().ref_linear_solver(self.h, b, _fid(f), _fid(prev), a, c)
    def diffuse(self, b, f, prev): self.lib().ref_diffuse(self.h, b, _fid(f), _fid(prev))
    def project(self): self.lib().ref_project(self.h)
    def advect(self, b, f, prev): self.lib().ref_advect(self.h, b, _fid(f), _fid(prev))
    def step(self): self.lib().ref_step(self.h)
    def run_iterations(self, n=1): self.lib().ref_run_iterations(self.h, n)

    def load_stl(self, path, scale=0.8, rot=(0.0, 0.0, 0.0), translate=(0.0, 0.0, 0.0)):
        self.lib().ref_load_stl(self.h, os.fsencode(path), scale, *rot, *translate)


class OracleSim(_Base):
    """The plain-C restatement (oracle/fluid_oracle.c)."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            L = C.CDLL(oracle_lib_path())
            L.fo_create.restype = _p
            L.fo_create.argtypes = [_i, _i, _i, _i, _f, _f, _f, _i, _i]
            L.fo_destroy.argtypes = [_p]
            L.fo_field.argtypes = [_p, _i]; L.fo_field.restype = C.POINTER(_f)
            L.fo_size.argtypes = [_p]; L.fo_size.restype = _i
            L.fo_set_order.argtypes = [_p, _i]
            L.fo_set_acc.argtypes = [_p, _i]
            L.fo_set_elide_dead.argtypes = [_p, _i]
            L.fo_add_obstacle.argtypes = [_p, _i, _i, _i]
            L.fo_add_density.argtypes = [_p, _i, _i, _i, _f]
            L.fo_set_velocity.argtypes = [_p, _i, _i, _i, _f, _f, _f]
            L.fo_set_bounds.argtypes = [_p, _i, _i]
            L.fo_linear_solver.argtypes = [_p, _i, _i, _i, _f, _f]
            L.fo_diffuse.argtypes = [_p, _i, _i, _i]
            L.fo_divergence.argtypes = [_p]
            L.fo_gradient.argtypes = [_p]
            L.fo_project.argtypes = [_p]
            L.fo_advect.argtypes = [_p, _i, _i, _i]
            L.fo_inflow_density.argtypes = [_p]
            L.fo_inflow_velocity.argtypes = [_p]
            L.fo_step.argtypes = [_p]
            L.fo_run_iteration.argtypes = [_p]
            L.fo_run_iterations.argtypes = [_p, _i]
            L.fo_diffuse_a.argtypes = [_p]; L.fo_diffuse_a.restype = _f
            L.fo_project_h.argtypes = [_p]; L.fo_project_h.restype = _f
            cls._lib = L
        return cls._lib

    def __init__(self, w, h, d, iter=1, speed=30, dt=0.05, diff=2.0e-5, visc=1.5e-5, acc=15,
                 order=ORDER_LEX, elide_dead=False):
        super().__init__(w, h, d)
        self.h = self.lib().fo_create(w, h, d, speed, dt, diff, visc, acc, order)
        self.acc = acc
        if elide_dead:
            self.lib().fo_set_elide_dead(self.h, 1)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib().fo_destroy(self.h)
            self.h = None

    def field(self, f):
        return self._view(self.lib().fo_field(self.h, _fid(f)))

    def set_order(self, order): self.lib().fo_set_order(self.h, order)

    def set_acc(self, acc):
        self.acc = acc
        self.lib().fo_set_acc(self.h, acc)

    def add_obstacle(self, x, y, z): self.lib().fo_add_obstacle(self.h, x, y, z)
    def add_density(self, x, y, z, a): self.lib().fo_add_density(self.h, x, y, z, a)
    def set_velocity(self, x, y, z, ax, ay, az): self.lib().fo_set_velocity(self.h, x, y, z, ax, ay, az)
    def set_bounds(self, b, f): self.lib().fo_set_bounds(self.h, b, _fid(f))
    def linear_solver(self, b, f, prev, a, c): self.lib().fo_linear_solver(self.h, b, _fid(f), _fid(prev), a, c)
    def diffuse(self, b, f, prev): self.lib().fo_diffuse(self.h, b, _fid(f), _fid(prev))
    def divergence(self): self.lib().fo_divergence(self.h)
    def gradient(self): self.lib().fo_gradient(self.h)
    def project(self): self.lib().fo_project(self.h)
    def advect(self, b, f, prev): self.lib().fo_advect(self.h, b, _fid(f), _fid(prev))
    def inflow_density(self): self.lib().fo_inflow_density(self.h)
    def inflow_velocity(self): self.lib().fo_inflow_velocity(self.h)
    def step(self): self.lib().fo_step(self.h)
    def run_iterations(self, n=1): self.lib().fo_run_iterations(self.h, n)
    def diffuse_a(self): return self.lib().fo_diffuse_a(self.h)
    def project_h(self): return self.lib().fo_project_h(self.h)


# ---- STL voxeliser checkers (object_loader.cpp:288-449) ---------------------------------------------
class _VoSetup(C.Structure):
    _fields_ = [("center", C.c_float * 3), ("max_radius", C.c_float), ("pmin", C.c_float * 3), ("pmax", C.c_float * 3),
                ("obj_size", C.c_float), ("resolution", C.c_float), ("coarse_res", C.c_float), ("grid_scale", C.c_float),
                ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int)]


def voxelize(tris, w, h, d, scale=0.8, rot=(0.0, 0.0, 0.0), translate=(0.0, 0.0, 0.0), mode=0):
    """oracle/voxel_oracle.c: (padded obs mask (d+2, h+2, w+2) float32, added count, setup dict)."""
    lib = OracleSim.lib()
    lib.vo_voxelize.restype = C.c_longlong
    lib.vo_voxelize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_float] * 7 + \
                               [C.c_int, C.c_void_p, C.POINTER(_VoSetup)]
    tris = np.ascontiguousarray(np.asarray(tris, dtype=np.float32).reshape(-1, 9))
    obs = np.zeros((d + 2, h + 2, w + 2), np.float32)
    st = _VoSetup()
    added = lib.vo_voxelize(tris.ctypes.data, len(tris), w, h, d, scale, *[float(r) for r in rot],
                            *[float(t) for t in translate], mode, obs.ctypes.data, C.byref(st))
    setup = {k: (list(getattr(st, k)) if k in ("center", "pmin", "pmax") else getattr(st, k)) for k, _ in _VoSetup._fields_}
    return obs, int(added), setup


def ref_voxelize(path, w, h, d, scale=0.8, rot=(0.0, 0.0, 0.0), translate=(0.0, 0.0, 0.0), threads=None):
    """The reference's own loadSTLIntoObstacles on a fresh Simulation: padded obs mask (not reproducible run to run)."""
    if threads:
        RefSim.set_threads(threads)
    sim = RefSim(w, h, d)
    sim.load_stl(path, scale, rot, translate)
    return sim.get("obs").copy()
