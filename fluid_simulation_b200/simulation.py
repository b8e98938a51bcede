"""Python mirror of the reference's `Simulation` class (simulation.h:11-91) over libfluid_cuda.

Same constructor arguments, same public members and method names (`run`, `step`, `addObstacle`,
`addDensity`, `setVelocity`); the eleven std::vector<float> fields live in HBM behind a
`fluid_handle*` and are reached with `get`/`set` in the reference layout, shape (d+2, h+2, w+2).
The per-phase private methods of the reference are exposed for parity tests (`setBounds`,
`linearSolver`, `diffuse`, `project`, `advect`) through `fluid_call`.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib

FIELDS = {
    "v_x": 0, "v_y": 1, "v_z": 2, "v_x_prev": 3, "v_y_prev": 4, "v_z_prev": 5,
    "dens": 6, "buffer": 7, "obs": 8, "pressure": 9, "divergence": 10,
}
FIELD_NAMES = list(FIELDS)

SOLVER_RB, SOLVER_JACOBI, SOLVER_LEX = 0x0, 0x1, 0x2
VOX_HASHED, VOX_FIXED, VOX_CELLS = 0, 1, 2          # fluid_voxelize sampling (include/fluid_cuda.h)
ELIDE_DEAD, NO_TILED, TILED_ALWAYS = 0x10, 0x20, 0x40

OP_SET_BOUNDS, OP_LIN_SOLVE, OP_DIFFUSE, OP_DIVERGENCE, OP_GRADIENT = 1, 2, 3, 4, 5
OP_PROJECT, OP_ADVECT, OP_INFLOW_DENS, OP_INFLOW_VEL, OP_STEP, OP_COPY, OP_ADVECT_VEL = 6, 7, 8, 9, 10, 11, 12

FRAME_FILES = ("data.bin", "obs.bin", "v_x.bin", "v_y.bin", "v_z.bin")   # simulation.cpp:56-60
FRAME_FIELDS = ("dens", "obs", "v_x", "v_y", "v_z")                       # simulation.cpp:143-147


class FluidError(RuntimeError):
    pass


def _fid(f) -> int:
    return FIELDS[f] if isinstance(f, str) else int(f)


class Simulation:
    """Drop-in for the reference class; all compute runs in libfluid_cuda on the current GPU."""

    def __init__(self, w, h, d, iter, speed=30, dt=0.05, diff=2.0e-5, visc=1.5e-5, acc=15,
                 flags=SOLVER_RB, rank=0, nranks=1):
        self._L = _lib.load()
        self.width, self.height, self.depth = int(w), int(h), int(d)
        self.iter, self.speed = int(iter), int(speed)
        self.dt, self.diff, self.visc = float(dt), float(diff), float(visc)
        self._acc = int(acc)
        self.flags = int(flags)
        self.rank, self.nranks = rank, nranks
        self.shape = (self.depth + 2, self.height + 2, self.width + 2)
        self.size = self.shape[0] * self.shape[1] * self.shape[2]             # simulation.cpp:35
        if nranks == 1:
            self._h = self._L.fluid_create(w, h, d, speed, dt, diff, visc, acc, flags)
        else:
            self._h = self._L.fluid_create_slab(w, h, d, speed, dt, diff, visc, acc, flags, rank, nranks)
        if not self._h:
            raise FluidError(self._L.fluid_last_error(None).decode())
        zlo, dl = C.c_int(), C.c_int()
        self._L.fluid_get_dims(self._h, None, None, None, C.byref(zlo), C.byref(dl))
        self.z_lo, self.d_local = zlo.value, dl.value
        self._frame_bufs = None
        self._frame_files = None

    # -- plumbing ---------------------------------------------------------------------------
    def _ck(self, rc):
        if rc != 0:
            raise FluidError(f"libfluid_cuda error {rc}: {self._L.fluid_last_error(self._h).decode()}")

    def close(self):
        if getattr(self, "_h", None):
            if self._frame_files:
                self.flush_frames()
                for f in self._frame_files:
                    f.close()
                self._frame_files = None
            self._L.fluid_destroy(self._h)
            self._h = None
        if getattr(self, "_frame_bufs", None):
            for _, ptr in self._frame_bufs:
                self._L.fluid_host_free(ptr)
            self._frame_bufs = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def acc(self):
        return self._acc

    @acc.setter
    def acc(self, v):
        self._acc = int(v)
        self._ck(self._L.fluid_set_acc(self._h, self._acc))

    # -- field access (replaces direct use of the std::vectors) ------------------------------
    def set(self, f, arr):
        a = np.ascontiguousarray(np.asarray(arr, dtype=np.float32).reshape(self.shape))
        if _fid(f) == FIELDS["obs"]:
            self._ck(self._L.fluid_upload_obstacles(self._h, a.ctypes.data))
        else:
            self._ck(self._L.fluid_upload_field(self._h, _fid(f), a.ctypes.data))

    def get(self, f) -> np.ndarray:
        out = np.zeros(self.shape, dtype=np.float32)
        self._ck(self._L.fluid_download_field(self._h, _fid(f), out.ctypes.data))
        return out

    def state(self) -> dict:
        return {n: self.get(n) for n in FIELD_NAMES}

    def load_state(self, st: dict):
        for n, a in st.items():
            self.set(n, a)

    # -- the reference's public mutators (simulation.cpp:155-178) ----------------------------
    def addObstacle(self, x, y, z):
        xyz = np.array([x, y, z], dtype=np.int32)
        self._ck(self._L.fluid_add_obstacles(self._h, 1, xyz.ctypes.data))

    def addObstacles(self, xyz):
        xyz = np.ascontiguousarray(np.asarray(xyz, dtype=np.int32).reshape(-1, 3))
        self._ck(self._L.fluid_add_obstacles(self._h, len(xyz), xyz.ctypes.data))

    # -- loadSTLIntoObstacles (object_loader.cpp:288-449) ------------------------------------
    def voxelize(self, tris, scale=0.8, rot=(0.0, 0.0, 0.0), translate=(0.0, 0.0, 0.0), sampling=VOX_HASHED) -> int:
        """Rotate / scale / place a triangle mesh ((n, 9) floats) and mark the cells inside it as obstacle,
        on the device.  Returns the reference's "Added N obstacle points" count."""
        tris = np.ascontiguousarray(np.asarray(tris, dtype=np.float32).reshape(-1, 9))
        added = C.c_longlong(0)
        self._ck(self._L.fluid_voxelize(self._h, tris.ctypes.data, len(tris), float(scale), float(rot[0]), float(rot[1]),
                                        float(rot[2]), float(translate[0]), float(translate[1]), float(translate[2]),
                                        int(sampling), C.byref(added)))
        return int(added.value)

    def load_stl(self, path, scale=0.8, rot=(0.0, 0.0, 0.0), translate=(0.0, 0.0, 0.0), sampling=VOX_HASHED) -> int:
        """The reference's loadSTLIntoObstacles(path, sim, scale, rot_x, rot_y, rot_z, tx, ty, tz); an unreadable or
        empty file is reported and skipped like the reference does (object_loader.cpp:292-296), returning 0."""
        from .stl import read_stl
        try:
            tris = read_stl(path)
        except OSError:
            tris = np.zeros((0, 9), np.float32)
        if len(tris) == 0:
            import sys
            print(f"Failed to load STL: {path}", file=sys.stderr)
            return 0
        return self.voxelize(tris, scale, rot, translate, sampling)

    def addDensity(self, x, y, z, amount):
        xyz = np.array([x, y, z], dtype=np.int32)
        am = np.array([amount], dtype=np.float32)
        self._ck(self._L.fluid_add_density(self._h, 1, xyz.ctypes.data, am.ctypes.data))

    def setVelocity(self, x, y, z, amount_x, amount_y, amount_z):
        xyz = np.array([x, y, z], dtype=np.int32)
        v = np.array([amount_x, amount_y, amount_z], dtype=np.float32)
        self._ck(self._L.fluid_set_velocity(self._h, 1, xyz.ctypes.data, v.ctypes.data))

    def pinned_array(self, shape, dtype):
        """numpy view of a pinned host allocation (freed with the Simulation)."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        ptr = self._L.fluid_host_alloc(max(n, 1))
        if not ptr:
            raise FluidError("pinned host allocation failed")
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(ptr)
        buf = (C.c_char * max(n, 1)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def stage_field(self, arr):
        """Pinned host copy of the planes of a padded field that this rank holds (its own planes and its halo
        planes): make it once, outside any timed region, and hand it to `set_staged` -- the H2D copy then runs
        from pinned memory (70 MB at 256^3: 1.3 ms instead of ~6 ms from pageable memory)."""
        from .slab import HALO
        a = np.asarray(arr, dtype=np.float32).reshape(self.shape)
        lo = 0 if self.rank == 0 else 1 - HALO
        hi = self.d_local + 1 if self.rank == self.nranks - 1 else self.d_local + HALO
        first = self.z_lo + lo - 1
        buf = self.pinned_array((hi - lo + 1,) + tuple(self.shape[1:]), np.float32)
        buf[...] = a[first:first + hi - lo + 1]
        return buf, first

    def set_staged(self, f, staged):
        """`set(f, arr)` from a `stage_field(arr)` buffer."""
        buf, first = staged
        ptr = buf.ctypes.data - first * self.shape[1] * self.shape[2] * 4     # the C ABI takes the base of a whole field
        if _fid(f) == FIELDS["obs"]:
            self._ck(self._L.fluid_upload_obstacles(self._h, ptr))
        else:
            self._ck(self._L.fluid_upload_field(self._h, _fid(f), ptr))

    def add_density_cells(self, xyz, amount):
        """addDensity for n cells at once; xyz int32 (n,3), amount float32 (n,) (host arrays)."""
        self._ck(self._L.fluid_add_density(self._h, len(xyz), xyz.ctypes.data, amount.ctypes.data))

    def set_velocity_cells(self, xyz, v):
        self._ck(self._L.fluid_set_velocity(self._h, len(xyz), xyz.ctypes.data, v.ctypes.data))

    # -- the reference's private phases, for parity tests (simulation.h:33-40) ---------------
    def _call(self, op, b=0, f=0, prev=0, a=0.0, c=0.0):
        self._ck(self._L.fluid_call(self._h, op, b, _fid(f), _fid(prev), a, c))

    def setBounds(self, b, f): self._call(OP_SET_BOUNDS, b, f)
    def linearSolver(self, b, f, prev, a, c): self._call(OP_LIN_SOLVE, b, f, prev, a, c)
    def diffuse(self, b, f, prev): self._call(OP_DIFFUSE, b, f, prev)
    def divergence(self): self._call(OP_DIVERGENCE)
    def gradient(self): self._call(OP_GRADIENT)
    def project(self): self._call(OP_PROJECT)
    def advect(self, b, f, prev): self._call(OP_ADVECT, b, f, prev)
    def advect_velocity(self): self._call(OP_ADVECT_VEL)      # simulation.cpp:125-127, one launch
    def inflow_density(self): self._call(OP_INFLOW_DENS)
    def inflow_velocity(self): self._call(OP_INFLOW_VEL)

    def step(self):
        """Simulation::step() (simulation.cpp:96-150) without the frame write."""
        self._call(OP_STEP)

    def run_iterations(self, n=1):
        """n iterations of run()'s loop body: inlet density, buffer = dens, step() (:65-71)."""
        self._ck(self._L.fluid_step(self._h, n))

    def sync(self):
        self._ck(self._L.fluid_sync(self._h))

    # -- statistics (simulation.cpp:73-90) ----------------------------------------------------
    def stats(self) -> dict:
        out = (C.c_double * 9)()
        self._ck(self._L.fluid_stats(self._h, out))
        keys = ("dens_sum", "dens_min", "dens_max", "vx_min", "vx_max", "vy_min", "vy_max", "vz_min", "vz_max")
        return dict(zip(keys, list(out)))

    def divergence_norm(self):
        out = (C.c_double * 2)()
        self._ck(self._L.fluid_divergence_norm(self._h, out))
        return out[0], out[1]

    # -- the offline viewers' arithmetic (gui.py:61-79, 265-300; GUI/utils.py:118-214) --------
    def render_slice(self, field="dens", z=None, vmin=None, vmax=None, obstacle_alpha=0.2) -> np.ndarray:
        """What gui.py's FluidViewer.update_image shows for plane `z` of `field`: an (h+2, w+2, 3) uint8
        image, colour-mapped with the viewers' density_cmap and darkened over the obstacle.  vmin / vmax
        default to the viewer's fixed ranges (gui.py:272-290)."""
        rng = {"dens": (0.0, 0.01), "v_x": (-10.0, 10.0), "v_y": (-1.0, 1.0), "v_z": (-1.0, 1.0)}
        name = field if isinstance(field, str) else FIELD_NAMES[int(field)]
        lo, hi = rng.get(name, (0.0, 1.0))
        vmin = lo if vmin is None else vmin
        vmax = hi if vmax is None else vmax
        z = (self.depth + 2) // 2 if z is None else int(z)
        out = np.zeros((self.height + 2, self.width + 2, 3), dtype=np.uint8)
        self._ck(self._L.fluid_render_slice(self._h, _fid(field), z, float(vmin), float(vmax), float(obstacle_alpha),
                                            out.ctypes.data))
        return out

    def streamlines(self, density=30, proximity=2, integration_steps=100, step_size=0.2, velocity_change_threshold=0.1):
        """GUI/utils.py generate_streamlines on the current velocity and obstacle fields (defaults =
        GUI/config.py:19-24).  Returns (lines, norm_speed): a list of (n_i, 3) float64 arrays of padded
        [x, y, z] positions in the reference's order, and per line the argument of its speed colour."""
        from .viewer import StreamlineParams
        sp = StreamlineParams(int(density), float(proximity), int(integration_steps), float(step_size),
                              float(velocity_change_threshold))
        half = sp.integration_steps // 2
        max_lines = max(1, sp.density * (sp.density // 2) ** 2)
        max_points = max_lines * (2 * half + 1)
        nl, npnt = C.c_int(0), C.c_longlong(0)
        # sizes first (no copy), then buffers that fit exactly
        rc = self._L.fluid_streamlines(self._h, C.byref(sp), 0, 0, C.byref(nl), C.byref(npnt), None, None, None)
        if rc != 0 and not (nl.value or npnt.value):
            self._ck(rc)
        if nl.value == 0:
            return [], np.zeros(0)
        assert nl.value <= max_lines and npnt.value <= max_points
        lens = np.zeros(nl.value, np.int32); norm = np.zeros(nl.value); pts = np.zeros((npnt.value, 3))
        self._ck(self._L.fluid_streamlines(self._h, C.byref(sp), nl.value, npnt.value, C.byref(nl), C.byref(npnt),
                                           lens.ctypes.data, norm.ctypes.data, pts.ctypes.data))
        return np.split(pts, np.cumsum(lens)[:-1]), norm

    def obstacle_mesh(self, level=0.5) -> dict:
        """GUI/utils.py:10-38 generate_obstacle_mesh on the resident obstacle field: marching cubes at `level`.
        Returns the viewer's dict: 'vertexes' (n, 3) float32 in padded [x, y, z] index coordinates, 'faces' (m, 3)
        int32, 'vertex_colors' (n, 4) opaque grey; empty arrays when there is no obstacle (utils.py:31-37)."""
        nv, nf = C.c_longlong(0), C.c_longlong(0)
        self._ck(self._L.fluid_obstacle_mesh(self._h, float(level), 0, 0, C.byref(nv), C.byref(nf), None, None))
        if nv.value == 0:
            return {"vertexes": np.array([]), "faces": np.array([]), "vertex_colors": np.array([])}
        verts = np.zeros((nv.value, 3), np.float32)
        faces = np.zeros((max(nf.value, 1), 3), np.int32)
        self._ck(self._L.fluid_obstacle_mesh(self._h, float(level), nv.value, faces.shape[0], C.byref(nv), C.byref(nf),
                                             verts.ctypes.data, faces.ctypes.data))
        colors = np.ones((nv.value, 4))
        colors[:, :3] = 0.5                                       # utils.py:20-24
        return {"vertexes": verts, "faces": faces[:nf.value], "vertex_colors": colors}

    def launch_count(self) -> int:
        return int(self._L.fluid_launch_count(self._h))

    def profile(self, on=True):
        self._ck(self._L.fluid_profile_enable(self._h, int(on)))

    def profile_read(self):
        """(ms inside linearSolver, field-sweeps, solver calls) since the last read."""
        out = (C.c_double * 3)()
        self._ck(self._L.fluid_profile_read(self._h, out))
        return out[0], out[1], out[2]

    def timer_start(self):
        self._ck(self._L.fluid_timer_start(self._h))

    def timer_stop(self) -> float:
        ms = C.c_float()
        self._ck(self._L.fluid_timer_stop(self._h, C.byref(ms)))
        return ms.value

    # -- frame dumps (simulation.cpp:56-60, 140-148) ------------------------------------------
    def _ensure_frame_bufs(self):
        """Five pinned host buffers for (dens, obs, v_x, v_y, v_z), each holding the padded planes this rank OWNS
        (the whole frame on one GPU; a z-slab's share of it otherwise, so that eight ranks of a 1024 x 512 x 512
        run pin 0.7 GB each instead of 5.4 GB)."""
        if self._frame_bufs is None:
            first, count = self._owned_planes()
            plane = self.shape[1] * self.shape[2]
            bufs = []
            for _ in range(5):
                ptr = self._L.fluid_host_alloc(count * plane * 4)
                if not ptr:
                    raise FluidError("pinned host allocation failed")
                arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float)), shape=(count * plane,))
                bufs.append((arr.reshape((count,) + self.shape[1:]), ptr))
            self._frame_bufs = bufs
        return self._frame_bufs

    def dump_frame_async(self, which=(True, True, True, True, True)):
        """Start the async D2H of (dens, obs, v_x, v_y, v_z) into pinned buffers; returns the views (this rank's
        owned planes, `_owned_planes()[0]` first)."""
        bufs = self._ensure_frame_bufs()
        first, _ = self._owned_planes()
        # the C ABI takes the base of a whole-frame buffer and writes this rank's planes at their offset in it
        skew = first * self.shape[1] * self.shape[2] * 4
        ptrs = (C.c_void_p * 5)(*[(ptr - skew) if w else None for (_, ptr), w in zip(bufs, which)])
        self._ck(self._L.fluid_dump_frame_async(self._h, ptrs))
        return [a for a, _ in bufs]

    def frame_wait(self):
        self._ck(self._L.fluid_frame_wait(self._h))

    def _owned_planes(self):
        """(first global padded plane, count) this rank owns in a frame: its interior planes plus the global
        ghost plane(s) of the end slabs -- the planes fluid_dump_frame_async fills."""
        first = 0 if self.rank == 0 else self.z_lo
        last = self.depth + 1 if self.rank == self.nranks - 1 else self.z_lo + self.d_local - 1
        return first, last - first + 1

    def open_frame_files(self, directory="data", dist=None):
        """Open the reference's five output files (simulation.cpp:56-60).  z-slabs (nranks > 1) share ONE set
        of files: rank 0 creates and truncates them, every rank then writes only the planes it owns at their
        offset inside each frame (`os.pwrite`), so the bytes equal the single-GPU / reference files.  That is a
        collective and needs the process group (`dist`); without it a multi-rank handle refuses."""
        if self.nranks > 1 and dist is None:
            raise FluidError("open_frame_files: a z-slab handle (nranks > 1) needs the process group (dist=...): "
                             "ranks share the frame files and must not truncate each other's frames")
        self._frame_dist = dist if self.nranks > 1 else None
        self._frames_written = 0
        if self._frame_dist is None or self.rank == 0:
            os.makedirs(directory, exist_ok=True)
            self._frame_files = [open(os.path.join(directory, n), "wb") for n in FRAME_FILES]
        if self._frame_dist is not None:
            self._frame_dist.barrier()                   # the files exist and are empty
            if self.rank != 0:
                self._frame_files = [open(os.path.join(directory, n), "r+b") for n in FRAME_FILES]

    def flush_frames(self):
        if self._frame_files and getattr(self, "_frame_pending", False):
            self.frame_wait()
            if getattr(self, "_frame_dist", None) is None:
                if self.nranks > 1:
                    raise FluidError("flush_frames: multi-rank frame files were not opened with dist=")
                for f, (arr, _) in zip(self._frame_files, self._frame_bufs):
                    f.write(arr.tobytes())
            else:
                first, count = self._owned_planes()
                plane_bytes = self.shape[1] * self.shape[2] * 4
                off = self._frames_written * self.size * 4 + first * plane_bytes
                for f, (arr, _) in zip(self._frame_files, self._frame_bufs):
                    os.pwrite(f.fileno(), arr.tobytes(), off)
            self._frames_written = getattr(self, "_frames_written", 0) + 1
            self._frame_pending = False

    # -- restart: a dumped frame is a complete state ----------------------------
    def write_sidecar(self, directory, frames, dump_every=1):
        """data/meta.json: what the reference's viewers keep in hand-synchronised constants
        (GUI/config.py:8-11, gui.py:32-34) and what a restart needs."""
        import json
        meta = {"width": self.width, "height": self.height, "depth": self.depth, "padded": [self.depth + 2, self.height + 2, self.width + 2],
                "frames": int(frames), "dump_every": int(dump_every), "dtype": "float32", "order": "z,y,x (C order, padded)",
                "files": list(FRAME_FILES), "speed": self.speed, "dt": self.dt, "diff": self.diff, "visc": self.visc, "acc": self.acc}
        with open(os.path.join(directory, "meta.json"), "w") as f:
            json.dump(meta, f, indent=1)

    def restart_from_frames(self, directory="data", frame=-1):
        """Load frame `frame` (negative = from the end) of data/{data,obs,v_x,v_y,v_z}.bin as the current
        state.  The other six fields are scratch that step() overwrites before reading (v_*_prev and buffer
        are copied from v / dens, pressure and divergence are recomputed, simulation.cpp:96-150), so the
        steps that follow reproduce the original run bit for bit.  Returns the frame index loaded."""
        npad = (self.depth + 2) * (self.height + 2) * (self.width + 2)
        sizes = [os.path.getsize(os.path.join(directory, n)) for n in FRAME_FILES]
        frames = min(sizes) // (4 * npad)
        if frames < 1 or any(sz % (4 * npad) for sz in sizes):
            raise FluidError(f"{directory}: frame files do not hold whole {self.width}x{self.height}x{self.depth} frames")
        k = frame if frame >= 0 else frames + frame
        if not 0 <= k < frames:
            raise FluidError(f"{directory}: frame {frame} out of range (0..{frames - 1})")
        for name, field in zip(FRAME_FILES, ("dens", "obs", "v_x", "v_y", "v_z")):
            a = np.fromfile(os.path.join(directory, name), dtype=np.float32, count=npad, offset=4 * npad * k)
            self.set(field, a.reshape(self.depth + 2, self.height + 2, self.width + 2))
        return k

    def run(self, directory="data", dump_every=1, verbose=True, dist=None):
        """Simulation::run() (simulation.cpp:49-91): `iter` steps, frames appended to data/*.bin,
        density sum every 100 steps and the final min/max statistics on stdout.  z-slab handles (nranks > 1)
        must pass the process group: frames go to shared files (see open_frame_files), statistics are reduced
        over the ranks, rank 0 prints and writes meta.json."""
        if self.nranks > 1 and dist is None:
            raise FluidError("run(): a z-slab handle (nranks > 1) needs the process group (dist=...); "
                             "per-rank runs would overwrite each other's frames and print per-slab statistics")
        lead = self.rank == 0
        verbose = verbose and lead

        def stats():
            st = self.stats()
            if self.nranks > 1:
                from . import multi_gpu
                st = multi_gpu.reduce_stats(dist, st)
            return st

        if verbose:
            print(f"starting 3-D simulation: {self.width}x{self.height}x{self.depth}  steps = {self.iter}")
        if directory is not None:
            self.open_frame_files(directory, dist)
        self._frame_pending = False
        for i in range(self.iter):
            self.run_iterations(1)
            if directory is not None and (i + 1) % dump_every == 0:
                self.flush_frames()              # write the previous frame while this step ran
                self.dump_frame_async()
                self._frame_pending = True
            if (i + 1) % 100 == 0 and i > 0:
                st = stats()                     # collective on z-slabs: every rank takes part
                if verbose:
                    print(f"step {i + 1}\n  density sum = {st['dens_sum']:g}")
        self.flush_frames()
        if directory is not None:
            for f in self._frame_files:
                f.flush()
            if self.nranks > 1:
                dist.barrier()                   # every rank's planes are in the files
            if lead:
                self.write_sidecar(directory, self.iter // dump_every, dump_every)
        st = stats()
        if verbose:
            print("\n--- statistics -------------------------------------------------")
            print(f"density  min = {st['dens_min']:g}\ndensity  max = {st['dens_max']:g}")
            for ax in "xyz":
                print(f"velocity {ax} min = {st['v' + ax + '_min']:g}\nvelocity {ax} max = {st['v' + ax + '_max']:g}")
            print("simulation finished")
        return st
