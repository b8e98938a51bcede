"""oracle/postproc.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes front-end of oracle/postproc_oracle.c (streamlines, slice image) and a NumPy restatement of the
colour table the reference's viewers build with matplotlib (gui.py:38-41, GUI/config.py:14-17).
matplotlib is not installed here, so the table follows matplotlib's published algorithm
(colors.LinearSegmentedColormap.from_list + colors._create_lookup_table, N = 256, gamma = 1) and is
PARITY UNPINNED; the streamline functions are pinned by tests/golden/postproc_*.npz, produced by the
reference's own GUI/utils.py (tests/golden/make_postproc_golden.py).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import oracle_lib_path

_lib = None

# CSS4 colours named by the viewers
_NAMED = {"white": "#FFFFFF", "lightgreen": "#90EE90", "green": "#008000", "deepskyblue": "#00BFFF",
          "blue": "#0000FF", "darkred": "#8B0000", "red": "#FF0000"}
DENSITY_CMAP_COLOURS = ["white", "lightgreen", "green", "deepskyblue", "blue", "darkred", "red"]


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(oracle_lib_path())
        L.oracle_streamlines.restype = C.c_int
        L.oracle_streamlines.argtypes = [C.c_void_p] * 4 + [C.c_int] * 4 + [C.c_double, C.c_int, C.c_double, C.c_double,
                                                                             C.c_int, C.c_longlong, C.c_void_p, C.c_void_p,
                                                                             C.c_void_p, C.POINTER(C.c_longlong)]
        L.oracle_render_slice.restype = None
        L.oracle_render_slice.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                          C.c_void_p, C.c_void_p]
        L.mo_marching_cubes.restype = C.c_int
        L.mo_marching_cubes.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.POINTER(C.c_longlong),
                                        C.POINTER(C.c_longlong), C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def create_lookup_table(n, data):
    """matplotlib.colors._create_lookup_table for gamma = 1; data = rows of (x, y0, y1)."""
    adata = np.array(data)
    x, y0, y1 = adata[:, 0], adata[:, 1], adata[:, 2]
    x = x * (n - 1)
    xind = (n - 1) * np.linspace(0, 1, n) ** 1.0
    ind = np.searchsorted(x, xind)[1:-1]
    distance = (xind[1:-1] - x[ind - 1]) / (x[ind] - x[ind - 1])
    lut = np.concatenate([[y1[0]], distance * (y0[ind] - y1[ind - 1]) + y1[ind - 1], [y0[-1]]])
    return np.clip(lut, 0.0, 1.0)


def density_cmap_lut(n=256) -> np.ndarray:
    """(n, 3) float64: LinearSegmentedColormap.from_list("density_cmap", DENSITY_CMAP_COLOURS)._lut[:n, :3]."""
    rgb = np.array([[int(_NAMED[c][i:i + 2], 16) / 255 for i in (1, 3, 5)] for c in DENSITY_CMAP_COLOURS])
    vals = np.linspace(0, 1, len(rgb))
    return np.stack([create_lookup_table(n, np.column_stack([vals, rgb[:, c], rgb[:, c]])) for c in range(3)], axis=1)


def streamlines(vx, vy, vz, obs, density=30, proximity=2, integration_steps=100, step_size=0.2, threshold=0.1):
    """generate_streamlines on (z, y, x)-ordered padded float32 volumes -> (list of (n, 3) arrays, norm_speed)."""
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (vx, vy, vz, obs)]
    nz, ny, nx = arrs[0].shape
    max_lines = max(1, density * (density // 2) ** 2)
    max_pts = max_lines * (2 * (integration_steps // 2) + 1)
    lens = np.zeros(max_lines, np.int32); norm = np.zeros(max_lines); pts = np.zeros((max_pts, 3))
    npts = C.c_longlong(0)
    n = lib().oracle_streamlines(*[a.ctypes.data for a in arrs], nx, ny, nz, density, float(proximity), integration_steps,
                                 float(step_size), float(threshold), max_lines, max_pts, lens.ctypes.data, norm.ctypes.data,
                                 pts.ctypes.data, C.byref(npts))
    assert n >= 0
    lens = lens[:n]
    return (np.split(pts[:npts.value], np.cumsum(lens)[:-1]) if n else []), norm[:n].copy()


def render_slice(field, obs, z, vmin, vmax, alpha=0.2) -> np.ndarray:
    """apply_cmap + overlay_obstacle (gui.py:61-79) on plane z of (z, y, x) volumes -> (ny, nx, 3) uint8."""
    f = np.ascontiguousarray(field[z], dtype=np.float32); o = np.ascontiguousarray(obs[z], dtype=np.float32)
    lut = np.ascontiguousarray(density_cmap_lut())
    out = np.zeros(f.shape + (3,), np.uint8)
    lib().oracle_render_slice(f.ctypes.data, o.ctypes.data, f.shape[1], f.shape[0], float(vmin), float(vmax), float(alpha),
                              lut.ctypes.data, out.ctypes.data)
    return out


def marching_cubes(obs, level=0.5):
    """oracle/mesh_oracle.c on a padded obstacle frame (nz, ny, nx): (verts (n, 3) float32 in [x, y, z] index
    coordinates, faces (m, 3) int32) -- what GUI/utils.py:17 asks skimage for (parity unpinned, see the C file)."""
    a = np.ascontiguousarray(obs, dtype=np.float32)
    nz, ny, nx = a.shape
    nv, nf = C.c_longlong(0), C.c_longlong(0)
    ptr = a.ctypes.data
    assert lib().mo_marching_cubes(ptr, nx, ny, nz, level, C.byref(nv), C.byref(nf), None, None) == 0
    verts = np.zeros((nv.value, 3), np.float32)
    faces = np.zeros((nf.value, 3), np.int32)
    assert lib().mo_marching_cubes(ptr, nx, ny, nz, level, C.byref(nv), C.byref(nf), verts.ctypes.data, faces.ctypes.data) == 0
    return verts, faces
