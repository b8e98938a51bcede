#!/usr/bin/env python
"""Golden vectors for the STL voxeliser (object_loader.cpp:288-449), made in THIS container where the reference
compiles: the mesh, the call arguments, the reference's own mask (loadSTLIntoObstacles, one thread so the
thread-id-seeded RNG and the visiting order are fixed) and the oracle's masks for the three sampling modes.
    python tests/golden/make_voxel_golden.py
The tests check (a) the oracle against its own committed masks (drift), (b) the oracle against the reference's
committed mask statistically, (c) on the GPU, the CUDA voxeliser against the oracle bit for bit."""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import oracle
from fluid_simulation_b200 import stl

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = {
    "box_40x32x24": (stl.box_mesh(), (40, 32, 24), 0.8, (0.0, 0.0, 0.0), (0.0, 0.0, 0.0)),
    "box_rot_48x48x40": (stl.box_mesh(), (48, 48, 40), 0.7, (20.0, 35.0, 50.0), (3.0, -2.0, 1.5)),
    "sphere_36x36x36": (stl.sphere_mesh(1.0, 8, 16), (36, 36, 36), 0.8, (0.0, 10.0, 0.0), (0.0, 0.0, 0.0)),
}
for name, (tris, (w, h, d), scale, rot, tr) in CASES.items():
    tmp = os.path.join(tempfile.mkdtemp(), name + ".stl")
    stl.write_stl_binary(tmp, tris)
    ref = oracle.ref_voxelize(tmp, w, h, d, scale, rot, tr, threads=1)
    out = {"tris": tris, "dims": np.array([w, h, d]), "scale": np.float32(scale), "rot": np.array(rot, np.float32),
           "translate": np.array(tr, np.float32), "ref_mask": np.packbits(ref > 0)}
    for mode in (0, 1, 2):
        m, added, setup = oracle.voxelize(tris, w, h, d, scale, rot, tr, mode)
        out[f"mask{mode}"] = np.packbits(m > 0); out[f"added{mode}"] = np.int64(added)
        out["lattice"] = np.array([setup["nx"], setup["ny"], setup["nz"]])
        out["resolution"] = np.float32(setup["resolution"])
    np.savez_compressed(os.path.join(HERE, f"vox_{name}.npz"), **out)
    print(name, "ref cells", int((ref > 0).sum()), {k: int(v) for k, v in out.items() if k.startswith("added")})
