"""Headless stand-in for the reference's viewers (gui.py, GUI/): no Qt, no OpenGL (unavailable on this
box) -- the arithmetic they do before drawing runs as CUDA kernels in libfluid_cuda (csrc/postproc.cuh)
and the results go to files:

    python -m fluid_simulation_b200.viewer data/ --field dens --z 33 --ppm slice.ppm --streamlines lines.npz

reads the frame files the solver wrote (data/{data,obs,v_x,v_y,v_z}.bin + meta.json), uploads one frame and
writes the z-slice image gui.py:265-300 would show (binary PPM) and the streamlines GUI/main_window.py:222-238
would hand to the GL widget (npz: points, line_len, norm_speed, rgba).
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

from . import _lib


class StreamlineParams(C.Structure):
    """struct fluid_streamline_params (include/fluid_cuda.h); defaults in GUI/config.py:19-24"""
    _fields_ = [("density", C.c_int), ("proximity", C.c_double), ("integration_steps", C.c_int),
                ("step_size", C.c_double), ("velocity_change_threshold", C.c_double)]


def density_cmap_lut() -> np.ndarray:
    """(256, 3) float64 table of the viewers' density_cmap (gui.py:38-41), from the library."""
    lut = np.zeros(768)
    rc = _lib.load().fluid_density_cmap(lut.ctypes.data_as(C.POINTER(C.c_double)))
    if rc != 0:
        raise RuntimeError(f"fluid_density_cmap failed ({rc})")
    return lut.reshape(256, 3)


def speed_colours(norm_speed) -> np.ndarray:
    """config.density_cmap(min(norm_speed, 1.0)) for every line (GUI/utils.py:208): (n, 4) float64 RGBA."""
    lut = density_cmap_lut()
    t = np.asarray(norm_speed, dtype=np.float64) * 256.0
    k = np.where(t == 256.0, 255, t.astype(np.int64)).clip(0, 255)
    return np.concatenate([lut[k], np.ones((len(k), 1))], axis=1)


def write_ppm(path, rgb):
    """Binary PPM (P6) of an (rows, cols, 3) uint8 image."""
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (rgb.shape[1], rgb.shape[0]))
        f.write(rgb.tobytes())


def main(argv=None):
    import argparse
    from .simulation import Simulation
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("directory", help="directory holding data.bin, obs.bin, v_x.bin, v_y.bin, v_z.bin and meta.json")
    ap.add_argument("--frame", type=int, default=-1, help="frame index (negative = from the end; GUI/ shows the last)")
    ap.add_argument("--field", default="dens", choices=["dens", "v_x", "v_y", "v_z"])
    ap.add_argument("--z", type=int, default=None, help="padded z plane (default: the middle one)")
    ap.add_argument("--ppm", default=None, help="write the colour-mapped slice here")
    ap.add_argument("--streamlines", default=None, help="write the streamlines here (.npz)")
    ap.add_argument("--density", type=int, default=30)
    ap.add_argument("--proximity", type=float, default=2)
    ap.add_argument("--steps", type=int, default=100)
    a = ap.parse_args(argv)
    with open(os.path.join(a.directory, "meta.json")) as f:
        meta = json.load(f)
    sim = Simulation(meta["width"], meta["height"], meta["depth"], 0)
    try:
        k = sim.restart_from_frames(a.directory, a.frame)
        print(f"frame {k} of {a.directory}: {sim.width}x{sim.height}x{sim.depth}")
        if a.ppm:
            write_ppm(a.ppm, sim.render_slice(a.field, a.z))
            print(f"wrote {a.ppm}")
        if a.streamlines:
            lines, norm = sim.streamlines(a.density, a.proximity, a.steps)
            pts = np.concatenate(lines) - 1 if lines else np.zeros((0, 3))     # main_window.py:236 shifts the origin
            np.savez(a.streamlines, points=pts, line_len=np.array([len(l) for l in lines], np.int32),
                     norm_speed=norm, rgba=speed_colours(norm))
            print(f"wrote {a.streamlines}: {len(lines)} streamlines, {len(pts)} points")
    finally:
        sim.close()


if __name__ == "__main__":
    main()
