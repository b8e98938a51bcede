"""ctypes binding of libfluid_cuda.so (include/fluid_cuda.h).

The library is the product; this module only declares its C ABI to Python.  There is no CPU
fallback: if the shared object is missing the import of the solver fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfluid_cuda.so")

_f, _i, _u, _p = C.c_float, C.c_int, C.c_uint, C.c_void_p
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)
_dp = C.POINTER(C.c_double)

# every symbol include/fluid_cuda.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "fluid_create": (_p, [_i, _i, _i, _i, _f, _f, _f, _i, _u]),
    "fluid_create_slab": (_p, [_i, _i, _i, _i, _f, _f, _f, _i, _u, _i, _i]),
    "fluid_destroy": (None, [_p]),
    "fluid_set_acc": (_i, [_p, _i]),
    "fluid_set_params": (_i, [_p, _i, _f, _f, _f, _i]),
    "fluid_get_dims": (_i, [_p, _ip, _ip, _ip, _ip, _ip]),
    "fluid_upload_field": (_i, [_p, _i, _p]),
    "fluid_download_field": (_i, [_p, _i, _p]),
    "fluid_upload_obstacles": (_i, [_p, _p]),
    "fluid_add_obstacles": (_i, [_p, _i, _p]),
    "fluid_voxelize": (_i, [_p, _p, _i, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                            _i, C.POINTER(C.c_longlong)]),
    "fluid_add_density": (_i, [_p, _i, _p, _p]),
    "fluid_set_velocity": (_i, [_p, _i, _p, _p]),
    "fluid_step": (_i, [_p, _i]),
    "fluid_call": (_i, [_p, _i, _i, _i, _i, _f, _f]),
    "fluid_dump_frame_async": (_i, [_p, C.POINTER(_p)]),
    "fluid_frame_wait": (_i, [_p]),
    "fluid_stats": (_i, [_p, _dp]),
    "fluid_divergence_norm": (_i, [_p, _dp]),
    "fluid_stats_static": (_i, [_p, _p, _p, _p, C.c_longlong, _p, _p]),
    "fluid_render_slice": (_i, [_p, _i, _i, C.c_double, C.c_double, C.c_double, _p]),
    "fluid_density_cmap": (_i, [_dp]),
    "fluid_streamlines": (_i, [_p, _p, _i, C.c_longlong, _ip, C.POINTER(C.c_longlong), _p, _p, _p]),
    "fluid_obstacle_mesh": (_i, [_p, C.c_double, C.c_longlong, C.c_longlong, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), _p, _p]),
    "fluid_sync": (_i, [_p]),
    "fluid_last_error": (C.c_char_p, [_p]),
    "fluid_host_alloc": (_p, [C.c_size_t]),
    "fluid_host_free": (None, [_p]),
    "fluid_timer_start": (_i, [_p]),
    "fluid_timer_stop": (_i, [_p, _fp]),
    "fluid_profile_enable": (_i, [_p, _i]),
    "fluid_profile_read": (_i, [_p, _dp]),
    "fluid_launch_count": (C.c_longlong, [_p]),
    "fluid_version": (C.c_char_p, []),
    "fluid_set_device": (_i, [_i]),
    "fluid_comm_unique_id": (_i, [C.c_char_p, C.c_char_p]),
    "fluid_comm_init": (_i, [_p, C.c_char_p, C.c_char_p]),
    "fluid_ipc_export": (_i, [_p, _i, C.c_char_p]),
    "fluid_ipc_import": (_i, [_p, _i, _i, C.c_char_p]),
    "fluid_barrier": (_i, [_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load libfluid_cuda.so and type every entry point.  Raises if the library is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(fluid_simulation_b200 has no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)          # AttributeError if the .so lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib
