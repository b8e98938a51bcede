"""csrc/recip_div.h on the host: the gradient's division-free quotient (simulation.cpp:330-334, x / h and
x / (2h)) must be the IEEE quotient bit for bit.  The header's own exhaustive sweep (all 2^23 significands of
one binade, what the library runs before the first launch) is run for the divisors of the BASELINE shapes, and
the guarded entry point is compared with the division on zeros, denormals, huge values and random floats."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = r'''
#include "recip_div.h"
#include <string.h>
extern "C" int rd_verify(float s) { return fl_recip_div_verify(s) ? 1 : 0; }
extern "C" long rd_mismatches(float s, const float* x, long n)
{
    const float r = 1.0f / s;
    long bad = 0;
    for (long i = 0; i < n; ++i) {
        const float a = fl_recip_div(x[i], s, r), b = x[i] / s;
        if (memcmp(&a, &b, 4) != 0 && !(a != a && b != b)) ++bad;
    }
    return bad;
}
'''


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("recip_div")
    src = d / "rd.cpp"
    src.write_text(SRC)
    so = d / "librd.so"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-I", os.path.join(ROOT, "fluid_simulation_b200", "csrc"),
                    str(src), "-o", str(so)], check=True)
    L = ctypes.CDLL(str(so))
    L.rd_verify.argtypes = [ctypes.c_float]
    L.rd_mismatches.argtypes = [ctypes.c_float, ctypes.c_void_p, ctypes.c_long]
    L.rd_mismatches.restype = ctypes.c_long
    return L


def _h(w, h, d):
    libm = ctypes.CDLL("libm.so.6")
    libm.cbrtf.restype = ctypes.c_float
    libm.cbrtf.argtypes = [ctypes.c_float]
    return np.float32(1.0) / np.float32(libm.cbrtf(float(np.float32(w * h * d))))      # simulation.cpp:292


SHAPES = [(256, 256, 512), (1024, 512, 512), (100, 100, 100), (72, 40, 33), (24, 16, 12), (254, 254, 254)]


@pytest.mark.parametrize("w,h,d", SHAPES)
def test_exhaustive_sweep_passes_for_baseline_divisors(lib, w, h, d):
    hh = _h(w, h, d)
    assert lib.rd_verify(hh) == 1 and lib.rd_verify(np.float32(2.0) * hh) == 1


def test_divisors_outside_the_range_keep_the_division(lib):
    assert lib.rd_verify(np.float32(2.0 ** -30)) == 0 and lib.rd_verify(np.float32(1024.0)) == 0
    assert lib.rd_verify(np.float32("nan")) == 0


@pytest.mark.parametrize("w,h,d", SHAPES[:4])
def test_guarded_quotient_equals_division_on_all_kinds_of_floats(lib, w, h, d):
    rng = np.random.default_rng(w + d)
    n = 1 << 21
    bits = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)                  # every exponent, both signs
    special = np.array([0x00000000, 0x80000000, 0x00000001, 0x807fffff, 0x00800000, 0x7f7fffff, 0xff7fffff,
                        0x7f800000, 0xff800000, 0x7fc00000, 0x21800000, 0x217fffff, 0x5d800000, 0x5d800001],
                       dtype=np.uint32)                                                       # +-0, denormals, extremes, the guard's edges
    x = np.ascontiguousarray(np.concatenate([bits, special]).view(np.float32))
    for s in (_h(w, h, d), np.float32(2.0) * _h(w, h, d)):
        assert lib.rd_mismatches(s, x.ctypes.data, x.size) == 0
