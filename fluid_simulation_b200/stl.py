"""STL reader with the reference's acceptance rules (object_loader.cpp:98-174) and small mesh generators
for tests and the bench.  The triangles go to the CUDA voxeliser as an (n, 9) float32 array
(v1, v2, v3 per row); normals and the 2-byte attribute are skipped as the reference skips them.

Rules kept from readSTL:
  * the file is binary unless its first line, stripped of blanks, starts with "solid" (:107-110);
  * binary: 80-byte header, uint32 count, 50 bytes per facet (:112-141);
  * ASCII: "vertex x y z" lines between "outer loop" and "endfacet"; a facet is kept only when exactly
    three vertices were read since its "outer loop"; a fourth vertex wraps the counter to 0 and the
    facet is dropped (:147-169);
  * nothing read -> empty list, which loadSTLIntoObstacles reports and returns on (:292-296).
"""
from __future__ import annotations

import struct

import numpy as np


def read_stl(path) -> np.ndarray:
    with open(path, "rb") as f:
        raw = f.read()
    first = raw.split(b"\n", 1)[0].decode("latin-1").strip(" \t\n\r")
    if not first.startswith("solid"):
        if len(raw) < 84:
            return np.zeros((0, 9), np.float32)
        (n,) = struct.unpack_from("<I", raw, 80)
        n = min(n, (len(raw) - 84) // 50)                 # a truncated file yields what is there
        rec = np.frombuffer(raw, dtype=np.uint8, count=n * 50, offset=84).reshape(n, 50)
        return np.ascontiguousarray(rec[:, 12:48]).view("<f4").reshape(n, 9).astype(np.float32)
    tris, tri, vi = [], [None, None, None], 0
    for line in raw.decode("latin-1").split("\n")[1:]:
        line = line.strip(" \t\n\r")
        if line == "outer loop":
            vi = 0
        elif line == "endfacet":
            if vi == 3:
                tris.append([c for v in tri for c in v])
        elif line.startswith("vertex"):
            parts = line[6:].split()
            try:
                xyz = [float(np.float32(p)) for p in parts[:3]]
            except ValueError:
                continue
            if len(xyz) == 3:
                if vi < 3:
                    tri[vi] = xyz
                vi = (vi + 1) % 4
    return np.asarray(tris, dtype=np.float32).reshape(-1, 9)


def write_stl_binary(path, tris) -> None:
    tris = np.asarray(tris, dtype="<f4").reshape(-1, 9)
    with open(path, "wb") as f:
        f.write(b"binary stl written by fluid_simulation_b200".ljust(80, b" "))
        f.write(struct.pack("<I", len(tris)))
        for t in tris:
            f.write(struct.pack("<3f", 0.0, 0.0, 0.0)); f.write(t.tobytes()); f.write(b"\0\0")


def write_stl_ascii(path, tris, name="mesh") -> None:
    tris = np.asarray(tris, dtype=np.float32).reshape(-1, 3, 3)
    with open(path, "w") as f:
        f.write(f"solid {name}\n")
        for t in tris:
            f.write("  facet normal 0 0 0\n    outer loop\n")
            for v in t:
                f.write(f"      vertex {float(v[0])!r} {float(v[1])!r} {float(v[2])!r}\n")
            f.write("    endloop\n  endfacet\n")
        f.write(f"endsolid {name}\n")


def box_mesh(lo=(-1.0, -0.5, -0.25), hi=(1.0, 0.5, 0.25)) -> np.ndarray:
    """Closed axis-aligned box, 12 triangles."""
    x0, y0, z0 = lo; x1, y1, z1 = hi
    v = np.array([[x0, y0, z0], [x1, y0, z0], [x1, y1, z0], [x0, y1, z0],
                  [x0, y0, z1], [x1, y0, z1], [x1, y1, z1], [x0, y1, z1]], np.float32)
    quads = [(0, 1, 2, 3), (4, 7, 6, 5), (0, 4, 5, 1), (1, 5, 6, 2), (2, 6, 7, 3), (3, 7, 4, 0)]
    tris = [[v[a], v[b], v[c]] for a, b, c, d in quads] + [[v[a], v[c], v[d]] for a, b, c, d in quads]
    return np.asarray(tris, np.float32).reshape(-1, 9)


def sphere_mesh(radius=1.0, n_lat=16, n_lon=32, center=(0.0, 0.0, 0.0)) -> np.ndarray:
    """Closed UV sphere, 2 * n_lon * (n_lat - 1) triangles."""
    th = np.linspace(0.0, np.pi, n_lat + 1)
    ph = np.linspace(0.0, 2 * np.pi, n_lon + 1)
    def pt(i, j):
        return [center[0] + radius * np.sin(th[i]) * np.cos(ph[j % n_lon]),
                center[1] + radius * np.sin(th[i]) * np.sin(ph[j % n_lon]),
                center[2] + radius * np.cos(th[i])]
    tris = []
    for i in range(n_lat):
        for j in range(n_lon):
            a, b, c, d = pt(i, j), pt(i + 1, j), pt(i + 1, j + 1), pt(i, j + 1)
            if i > 0:
                tris.append([a, b, d])
            if i < n_lat - 1:
                tris.append([b, c, d])
    return np.asarray(tris, np.float32).reshape(-1, 9)
