"""CPU tests of the drop-in boundary: libfluid_cuda.so loads and exports every symbol that
include/fluid_cuda.h declares; without a GPU the entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from fluid_simulation_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "fluid_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fluid_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert header_symbols() == sorted(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    for name in header_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.fluid_version()


def test_library_is_sm100a_only():
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", _lib.LIB_PATH],
                         capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs


def test_hot_path_kernels_are_fma_free_and_use_tma():
    """--fmad=false is the parity contract (no contraction of a*b+c): the kernels of the per-step hot path must
    hold no FFMA at all.  The FFMAs that do exist in the library live in kernels off that path, where they are
    library sequences for IEEE division / cbrtf / double arithmetic or integer-to-float index math, not
    contracted field arithmetic; the test lists where they are so that a new one on the hot path is caught.
    The sweep kernels must show the TMA and mbarrier instructions (UTMALDG, SYNCS)."""
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    per, name = {}, None
    for line in out.stdout.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1); per[name] = {"FFMA": 0, "UTMALDG": 0, "SYNCS": 0}
        elif name:
            for k in per[name]:
                if re.search(r"\b" + k + r"\b", line):
                    per[name][k] += 1
    hot = [n for n in per if re.search(r"k_sweep_|k_divergence|k_advect|k_inflow|k_set_bounds|k_copy_edges|k_lex_|k_pack|k_unpack|k_build", n)]
    assert hot and all(per[n]["FFMA"] == 0 for n in hot), {n: per[n]["FFMA"] for n in hot if per[n]["FFMA"]}
    # k_gradient<FL_GRAD_DIV> divides by h / 2h (IEEE division sequences contain FFMA by construction) and
    # <FL_GRAD_RECIP> holds the two explicit FMAs of csrc/recip_div.h's quotient; the power-of-two form the
    # benchmark grids take multiplies and must be clean
    grad = [n for n in per if "k_gradient" in n]
    assert any(per[n]["FFMA"] == 0 for n in grad), {n: per[n]["FFMA"] for n in grad}
    sweeps = [n for n in per if "k_sweep_stream" in n or "k_sweep_pair" in n]
    assert sweeps and all(per[n]["UTMALDG"] > 0 and per[n]["SYNCS"] > 0 for n in sweeps)


def _has_gpu():
    try:
        cuda = C.CDLL("libcudart.so")
    except OSError:
        return os.path.exists("/dev/nvidia0")
    n = C.c_int(0)
    return cuda.cudaGetDeviceCount(C.byref(n)) == 0 and n.value > 0


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="GPU present")
def test_no_cpu_fallback_without_gpu():
    from fluid_simulation_b200 import Simulation, FluidError
    with pytest.raises(FluidError, match="no CUDA device|no CPU fallback"):
        Simulation(8, 8, 8, 1)


def test_bad_arguments_are_rejected_without_device_work():
    lib = _lib.load()
    assert not lib.fluid_create(0, 8, 8, 30, 0.05, 2e-5, 1.5e-5, 15, 0)
    assert b"bad arguments" in lib.fluid_last_error(None)
    assert lib.fluid_sync(None) == -1
    assert lib.fluid_launch_count(None) == 0


def test_product_never_touches_the_oracle():
    """The product path must not import, link or load anything under oracle/."""
    pkg = os.path.join(ROOT, "fluid_simulation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in txt.replace("the CPU oracle", ""), os.path.join(dirpath, f)
    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "fluid_ref" not in out


def test_streamline_params_layout_matches_ctypes(tmp_path):
    """struct fluid_streamline_params crosses the ABI by pointer: the C layout (include/fluid_cuda.h) and the
    ctypes mirror (fluid_simulation_b200/viewer.py) must agree on size and every offset."""
    from fluid_simulation_b200.viewer import StreamlineParams
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "fluid_cuda.h"\nint main(void){'
                   'printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(fluid_streamline_params), offsetof(fluid_streamline_params, density),'
                   'offsetof(fluid_streamline_params, proximity), offsetof(fluid_streamline_params, integration_steps),'
                   'offsetof(fluid_streamline_params, step_size), offsetof(fluid_streamline_params, velocity_change_threshold));return 0;}')
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(StreamlineParams)] + [getattr(StreamlineParams, n).offset for n in
                                           ("density", "proximity", "integration_steps", "step_size", "velocity_change_threshold")]
    assert got == want


def test_header_compiles_as_plain_c(tmp_path):
    """The boundary is a C ABI: the header must be usable from a C translation unit (no C++ in the signatures)."""
    src = tmp_path / "h.c"
    src.write_text('#include "fluid_cuda.h"\nint main(void){ return FLUID_NFIELDS == 11 ? 0 : 1; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                    "-o", str(tmp_path / "h.o")], check=True)
