"""fluid_stats_static: the allocation-free variant of fluid_stats (include/fluid_cuda.h) against the existing
entry point.  Collected LAST on purpose.  Both go through the C ABI; torch only
provides the caller's device buffers -- everything is allocated BEFORE the calls under test, which allocate
nothing themselves (the per-block partials live in a file-scope device array of the library).

Tolerance: as tests/test_gpu_parity.py uses for fluid_stats -- the density sum to rtol 1e-12 (double
accumulation in a different order), the eight minima / maxima exactly."""
import ctypes as C

import numpy as np
import pytest


def _fields(w, h, d, seed):
    rng = np.random.default_rng(seed)
    shape = (d + 2, h + 2, w + 2)
    return {"dens": rng.random(shape, dtype=np.float32) * 3.0,
            "v_x": rng.standard_normal(shape, dtype=np.float32),
            "v_y": rng.standard_normal(shape, dtype=np.float32) * 0.5,
            "v_z": rng.standard_normal(shape, dtype=np.float32) * 2.0}


def _existing(L, w, h, d, fields):
    """fluid_stats on a handle holding the same padded fields."""
    from fluid_simulation_b200 import Simulation
    sim = Simulation(w, h, d, 1, acc=2)
    for name, a in fields.items():
        sim.set(name, a)
    out = (C.c_double * 9)()
    assert L.fluid_stats(sim._h, out) == 0, L.fluid_last_error(sim._h)
    return np.array(out[:], dtype=np.float64)


def _check(got, want, tag):
    assert np.isclose(got[0], want[0], rtol=1e-12), f"{tag}: sum(dens) {got[0]!r} vs {want[0]!r}"
    assert np.array_equal(got[1:], want[1:]), f"{tag}: minima / maxima {got[1:]} vs {want[1:]}"


@pytest.mark.gpu
def test_stats_static_matches_fluid_stats_on_two_calls_in_a_row():
    import torch
    from fluid_simulation_b200 import _lib
    L = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    # two volumes of different size and content: the second call must not see what the first left in the
    # shared array (the first uses 263 partials, the second 2)
    cases = [((254, 254, 254), 11), ((70, 40, 33), 12)]
    bufs = []
    for (w, h, d), seed in cases:
        f = _fields(w, h, d, seed)
        dev_f = {k: torch.from_numpy(v).to(dev).contiguous() for k, v in f.items()}
        bufs.append((f, dev_f, torch.full((9,), float("nan"), dtype=torch.float64, device=dev)))
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    for (f, dev_f, out), ((w, h, d), _) in zip(bufs, cases):          # back to back on one stream, no sync between
        n = (w + 2) * (h + 2) * (d + 2)
        rc = L.fluid_stats_static(dev_f["dens"].data_ptr(), dev_f["v_x"].data_ptr(), dev_f["v_y"].data_ptr(),
                                  dev_f["v_z"].data_ptr(), n, out.data_ptr(), stream.cuda_stream)
        assert rc == 0, L.fluid_last_error(None)
    stream.synchronize()
    for (f, dev_f, out), ((w, h, d), _) in zip(bufs, cases):
        got = out.cpu().numpy()
        _check(got, _existing(L, w, h, d, f), f"{w}x{h}x{d}")
        # and against numpy directly
        assert np.isclose(got[0], f["dens"].astype(np.float64).sum(), rtol=1e-12)
        assert got[3] == f["v_x"].min() and got[8] == f["v_z"].max()
    # the same volume again (third call): identical result, bit for bit
    f, dev_f, out = bufs[0]
    first = out.cpu().numpy().copy()
    out.fill_(float("nan"))
    (w, h, d), _ = cases[0]
    assert L.fluid_stats_static(dev_f["dens"].data_ptr(), dev_f["v_x"].data_ptr(), dev_f["v_y"].data_ptr(),
                                dev_f["v_z"].data_ptr(), (w + 2) * (h + 2) * (d + 2), out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), first)


@pytest.mark.gpu
def test_stats_static_rejects_sizes_beyond_the_static_scratch():
    import torch
    from fluid_simulation_b200 import _lib
    L = _lib.load()
    t = torch.zeros(16, dtype=torch.float32, device="cuda:0")
    out = torch.zeros(9, dtype=torch.float64, device="cuda:0")
    for n in ((1 << 28) + 1, 0, -5):
        rc = L.fluid_stats_static(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), n, out.data_ptr(), None)
        assert rc == -1                                       # FLUID_E_ARG, nothing launched
        assert b"n_cells" in L.fluid_last_error(None)
    assert L.fluid_stats_static(None, t.data_ptr(), t.data_ptr(), t.data_ptr(), 16, out.data_ptr(), None) == -1
    torch.cuda.synchronize()


def test_stats_static_is_declared_and_exported():
    """CPU: the header declares it, the library exports it, and its source calls no allocating routine."""
    import os
    import re
    from fluid_simulation_b200 import _lib
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    assert "fluid_stats_static" in open(os.path.join(root, "include", "fluid_cuda.h")).read()
    assert hasattr(_lib.load(), "fluid_stats_static")
    src = open(os.path.join(root, "fluid_simulation_b200", "csrc", "fluid_cuda.cu")).read()
    body = src[src.index("// ---- fluid_stats_static"):]
    assert "cudaGetSymbolAddress" in body
    body = re.sub(r"//[^\n]*", "", body)                       # the comments may name what the code must not call
    assert not re.search(r"cudaMalloc|cudaFree|cudaHostAlloc|\bnew\b|malloc\(|std::vector", body)
