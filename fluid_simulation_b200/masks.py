"""Analytic obstacle masks in the reference layout (padded float32, 1.0 = solid; simulation.h:22-23).

The reference's STL voxeliser is not reproducible run to run (thread-id seeded RNG,
object_loader.cpp:399), so benchmark and parity inputs use these deterministic masks and share
the bytes between the CPU reference and the GPU.
"""
from __future__ import annotations

import numpy as np


def _grid(w, h, d):
    z, y, x = np.meshgrid(np.arange(d + 2, dtype=np.float32), np.arange(h + 2, dtype=np.float32),
                          np.arange(w + 2, dtype=np.float32), indexing="ij")
    interior = (x >= 1) & (x <= w) & (y >= 1) & (y <= h) & (z >= 1) & (z <= d)
    return x, y, z, interior


def empty(w, h, d) -> np.ndarray:
    return np.zeros((d + 2, h + 2, w + 2), dtype=np.float32)


def sphere(w, h, d, centre=(0.35, 0.5, 0.5), radius=0.15) -> np.ndarray:
    """dist((x,y,z), (0.35W, 0.5H, 0.5D)) <= 0.15H, squared-distance test in float32."""
    x, y, z, interior = _grid(w, h, d)
    cx, cy, cz = np.float32(centre[0] * w), np.float32(centre[1] * h), np.float32(centre[2] * d)
    r = np.float32(radius * h)
    m = ((x - cx) ** 2 + (y - cy) ** 2 + (z - cz) ** 2 <= r * r) & interior
    return m.astype(np.float32)


def box(w, h, d, lo=(0.3, 0.35, 0.35), hi=(0.45, 0.65, 0.65)) -> np.ndarray:
    x, y, z, interior = _grid(w, h, d)
    m = ((x >= lo[0] * w) & (x <= hi[0] * w) & (y >= lo[1] * h) & (y <= hi[1] * h) &
         (z >= lo[2] * d) & (z <= hi[2] * d) & interior)
    return m.astype(np.float32)


def airfoil(w, h, d, chord=0.35, x0=0.25, span=(0.2, 0.8), thickness=0.12) -> np.ndarray:
    """Extruded NACA-00xx section (chord along x, thickness along y, span along z) -- the
    'car/airfoil' stand-in for BASELINE.json configs 3-5 (no mesh ships with the reference)."""
    x, y, z, interior = _grid(w, h, d)
    c = np.float32(chord * w)
    xi = np.clip((x - np.float32(x0 * w)) / c, 0.0, 1.0)
    yt = 5.0 * thickness * c * (0.2969 * np.sqrt(xi) - 0.1260 * xi - 0.3516 * xi ** 2 +
                                0.2843 * xi ** 3 - 0.1015 * xi ** 4)
    inside = (x >= x0 * w) & (x <= x0 * w + c) & (np.abs(y - np.float32(0.5 * h)) <= yt)
    m = inside & (z >= span[0] * d) & (z <= span[1] * d) & interior
    return m.astype(np.float32)


def airfoil_planes(w, h, d, z_first, z_count, **kw) -> np.ndarray:
    """The airfoil mask as a whole-tunnel array of which only the padded planes z_first .. z_first+z_count-1 are
    filled (the planes a z-slab holds): the section is the same in every plane of the span, so a 1024^3 mask costs
    one 2-D section and the planes asked for instead of 3 x 4.3 GB of coordinate grids per rank.  Untouched
    planes are untouched pages of a zero allocation."""
    sec = airfoil(w, h, 1, span=(0.0, 2.0), **kw)[1]                    # one interior plane: the (y, x) section
    span = kw.get("span", (0.2, 0.8))
    out = np.zeros((d + 2, h + 2, w + 2), dtype=np.float32)
    for z in range(max(z_first, 1), min(z_first + z_count, d + 1)):
        if np.float32(z) >= span[0] * d and np.float32(z) <= span[1] * d:
            out[z] = sec
    return out


def random_state(w, h, d, seed=0, obs=None) -> dict:
    """Seeded non-trivial state for per-function parity tests: v ~ U(-1,1), dens ~ U(0,0.01),
    every padded cell (ghosts, edges and corners included) filled."""
    rng = np.random.default_rng(seed)
    shape = (d + 2, h + 2, w + 2)
    st = {}
    for n in ("v_x", "v_y", "v_z", "v_x_prev", "v_y_prev", "v_z_prev", "pressure", "divergence"):
        st[n] = rng.uniform(-1.0, 1.0, shape).astype(np.float32)
    for n in ("dens", "buffer"):
        st[n] = rng.uniform(0.0, 0.01, shape).astype(np.float32)
    st["obs"] = sphere(w, h, d) if obs is None else np.asarray(obs, dtype=np.float32).reshape(shape)
    return st
