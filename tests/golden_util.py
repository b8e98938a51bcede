"""Load tests/golden/*.npz (made by tests/golden/make_golden.py from the reference object code)."""
import ast
import glob
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_cases():
    # solver goldens only; vox_*.npz belong to tests/test_voxelize.py, pp_*.npz to tests/test_postproc.py
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz"))
                  if not os.path.basename(p).startswith(("vox_", "pp_")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = ast.literal_eval(str(z["meta"]))
    inputs = {k[3:]: z[k] for k in z.files if k.startswith("in_")}
    outputs = {k[4:]: z[k] for k in z.files if k.startswith("out_")}
    return meta, inputs, outputs


def replay(sim, meta, api):
    """Run the golden case `meta` on `sim`; `api` maps generic op names to sim's methods."""
    case = meta.get("case")
    if case is None:
        api["run"](meta["steps"])
    elif case == "bounds":
        for b, f in ((0, "dens"), (1, "v_x"), (2, "v_y"), (3, "v_z")):
            api["bounds"](b, f)
    elif case == "linsolve":
        api["linsolve"](2, "v_y", "v_y_prev", 0.37, 3.22)
    elif case == "diffuse":
        api["diffuse"](1, "v_x", "v_x_prev")
    elif case == "project":
        api["project"]()
    elif case == "advect":
        for b, (f, p) in enumerate((("dens", "buffer"), ("v_x", "v_x_prev"), ("v_y", "v_y_prev"),
                                    ("v_z", "v_z_prev"))):
            api["advect"](b, f, p)
    elif case == "step":
        api["step"]()
    else:
        raise ValueError(case)
