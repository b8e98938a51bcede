// tiled.cuh -- TMA-staged shared-memory sweep kernels (placeholder until the tiled path lands:
// fl_tiled_solve reports "not handled" and the global-memory sweeps of kernels_stencil.cuh run).
#pragma once
#include "handle.cuh"

inline int fl_tiled_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                          float a, float c, bool* handled)
{
    (void)h; (void)nb; (void)b; (void)fid; (void)pid; (void)a; (void)c;
    *handled = false;
    return 0;
}

inline void fl_tiled_release(fluid_handle* h) { (void)h; }
