"""Generate tests/golden/*.npz from the UNMODIFIED reference object code (oracle/_ref).

Run in the build container (needs /root/reference to have been compiled by oracle/Makefile):
    python tests/golden/make_golden.py
Each file holds the seeded input state and the reference's output for one case, small enough to
commit.  The GPU box has no /root/reference; tests there check the oracle and the CUDA path against
these vectors (and against the prebuilt oracle/_ref when it travelled).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import RefSim, FIELD_NAMES  # noqa: E402
from fluid_simulation_b200 import masks  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def save(name, meta, inputs, outputs):
    d = {f"in_{k}": v for k, v in inputs.items()}
    d.update({f"out_{k}": v for k, v in outputs.items()})
    d["meta"] = np.array(repr(meta))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, {k: v for k, v in meta.items()})


def per_function(w, h, d, acc, seed):
    st = masks.random_state(w, h, d, seed=seed)
    for n in ("v_x", "v_y", "v_z", "v_x_prev", "v_y_prev", "v_z_prev"):
        st[n] = st[n] * np.float32(8.0)
    cases = {
        "bounds": lambda r: [r.set_bounds(b, f) for b, f in ((0, "dens"), (1, "v_x"), (2, "v_y"), (3, "v_z"))],
        "linsolve": lambda r: r.linear_solver(2, "v_y", "v_y_prev", 0.37, 3.22),
        "diffuse": lambda r: r.diffuse(1, "v_x", "v_x_prev"),
        "project": lambda r: r.project(),
        "advect": lambda r: [r.advect(b, f, p) for b, (f, p) in enumerate(
            (("dens", "buffer"), ("v_x", "v_x_prev"), ("v_y", "v_y_prev"), ("v_z", "v_z_prev")))],
        "step": lambda r: r.step(),
    }
    for cname, fn in cases.items():
        r = RefSim(w, h, d, acc=acc, threads=1)
        r.load_state(st)
        fn(r)
        save(f"fn_{cname}_{w}x{h}x{d}", dict(w=w, h=h, d=d, acc=acc, case=cname, seed=seed), st, r.state())


def from_rest(w, h, d, acc, steps, mask):
    obs = getattr(masks, mask)(w, h, d)
    r = RefSim(w, h, d, acc=acc, threads=1)
    r.set("obs", obs)
    r.run_iterations(steps)
    out = r.state()
    save(f"run_{mask}_{w}x{h}x{d}_acc{acc}_s{steps}", dict(w=w, h=h, d=d, acc=acc, steps=steps, mask=mask),
         {"obs": obs}, {k: out[k] for k in ("v_x", "v_y", "v_z", "dens", "pressure", "divergence")})


if __name__ == "__main__":
    per_function(12, 9, 7, 4, seed=11)
    from_rest(24, 16, 12, 7, 5, "sphere")
    from_rest(32, 16, 16, 10, 3, "empty")
