#!/usr/bin/env python
"""Golden vectors for the viewers' streamline arithmetic (GUI/utils.py:40-214), made in THIS container by the
reference's own Python: GUI/utils.py is imported from /root/reference unmodified, with the two modules it cannot
import here stubbed -- `skimage` (marching cubes only, unused by the streamline path) and `config`
(GUI/config.py needs matplotlib; the stub carries the same constants, and its density_cmap returns its argument
so the colour-map ARGUMENT of every line is recorded).
    python tests/golden/make_postproc_golden.py
Fields are seeded and smooth (a uniform stream, a swirl and two low modes around a sphere / box obstacle);
volumes are stored in the frame-file order (z, y, x), padded, float32."""
import os, sys, types
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_GUI = "/root/reference/GUI"


def fields(w, h, d, seed, box=False):
    rng = np.random.default_rng(seed)
    z, y, x = np.meshgrid(np.arange(d + 2), np.arange(h + 2), np.arange(w + 2), indexing="ij")
    cx, cy, cz = 0.4 * w, 0.5 * h + 0.5, 0.5 * d + 0.5
    if box:
        obs = (abs(x - cx) <= 0.12 * w) & (abs(y - cy) <= 0.2 * h) & (abs(z - cz) <= 0.22 * d)
    else:
        obs = (x - cx) ** 2 + (y - cy) ** 2 + (z - cz) ** 2 <= (0.22 * min(h, d)) ** 2
    ph = rng.uniform(0, 2 * np.pi, 6)
    r2 = (y - cy) ** 2 + (z - cz) ** 2 + 4.0
    vx = 3.0 + 1.5 * np.sin(2 * np.pi * y / h + ph[0]) * np.cos(2 * np.pi * z / d + ph[1])
    vy = -2.5 * (z - cz) / np.sqrt(r2) + 1.2 * np.sin(2 * np.pi * x / w * 2 + ph[2])
    vz = 2.5 * (y - cy) / np.sqrt(r2) + 1.2 * np.cos(2 * np.pi * x / w * 2 + ph[3]) * np.sin(2 * np.pi * y / h + ph[4])
    out = [a.astype(np.float32) for a in (vx, vy, vz)]
    for a in out:
        a[obs] = 0.0
    # a stagnant pocket (speed < 1e-6 stops a line, utils.py:92) behind the obstacle
    pocket = (abs(x - (cx + 0.3 * w)) <= 1) & (abs(y - cy) <= 1) & (abs(z - cz) <= 1)
    for a in out:
        a[pocket] = 0.0
    return out + [obs.astype(np.float32)]


CASES = {
    # name: (w, h, d, seed, box, density, proximity, steps, step_size, threshold)
    "sphere_28x18x16": (28, 18, 16, 11, False, 12, 2, 60, 0.2, 0.1),
    "box_24x20x14": (24, 20, 14, 5, True, 10, 30, 41, 0.35, 0.02),
    "defaults_40x24x24": (40, 24, 24, 3, False, 30, 2, 100, 0.2, 0.1),
}


def run_reference(vx, vy, vz, obs, density, proximity, steps, step_size, threshold):
    nz, ny, nx = vx.shape
    cfg = types.ModuleType("config")
    cfg.width, cfg.height, cfg.depth = nx, ny, nz
    cfg.STREAMLINE_DENSITY, cfg.STREAMLINE_PROXIMITY = density, proximity
    cfg.INTEGRATION_STEPS, cfg.INTEGRATION_STEP_SIZE, cfg.VELOCITY_CHANGE_THRESHOLD = steps, step_size, threshold
    cfg.density_cmap = lambda t: (float(t), 0.0, 0.0, 1.0)
    sk = types.ModuleType("skimage"); sk.measure = types.ModuleType("skimage.measure")
    saved = {k: sys.modules.get(k) for k in ("config", "skimage", "skimage.measure", "utils")}
    sys.modules.update({"config": cfg, "skimage": sk, "skimage.measure": sk.measure})
    sys.modules.pop("utils", None)
    sys.path.insert(0, REF_GUI)
    try:
        import utils                                    # /root/reference/GUI/utils.py, unmodified
        t = lambda a: np.transpose(a, (2, 1, 0))        # GUI/main_window.py:227-230
        lines, colours = utils.generate_streamlines(vx=t(vx), vy=t(vy), vz=t(vz), obs_data=t(obs), max_length=steps)
        # the sampler alone, at scattered positions incl. the clip range (utils.py:43-45)
        rng = np.random.default_rng(99)
        pos = rng.uniform(-1.5, 1.5 + max(nx, ny, nz), (200, 3))
        samp = np.array([utils._interpolate_scalar(t(vx), *p) for p in pos])
    finally:
        sys.path.remove(REF_GUI)
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return lines, np.array([c[0] for c in colours]), pos, samp


if __name__ == "__main__":
    for name, (w, h, d, seed, box, density, prox, steps, step_size, thr) in CASES.items():
        vx, vy, vz, obs = fields(w, h, d, seed, box)
        lines, norm, pos, samp = run_reference(vx, vy, vz, obs, density, prox, steps, step_size, thr)
        np.savez_compressed(os.path.join(HERE, f"pp_{name}.npz"), vx=vx, vy=vy, vz=vz, obs=obs,
                            params=np.array([density, prox, steps, step_size, thr], np.float64),
                            line_len=np.array([len(l) for l in lines], np.int32), norm_speed=norm,
                            points=np.concatenate(lines) if lines else np.zeros((0, 3)), sample_pos=pos, sample_val=samp)
        print(name, len(lines), "lines", sum(len(l) for l in lines), "points; numpy", np.__version__)
