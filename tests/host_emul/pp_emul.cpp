// tests/host_emul/pp_emul.cpp -- TEST HARNESS.  Compiles csrc/postproc_core.h (the statements the CUDA kernel
// k_pp_trace executes per seed) with g++ and runs them on the host, so that the CPU suite can compare the
// product's streamline source with the checker bit for bit where no GPU is at hand.  Not linked into the product.
#include "../../fluid_simulation_b200/csrc/postproc_core.h"
#include <vector>
#include <cstring>

extern "C" int emul_streamlines(const float* vx, const float* vy, const float* vz, const float* obs, int nx, int ny, int nz,
                                int density, int integration_steps, double step_size, double threshold,
                                const double* bmin, const double* bmax, double denom,
                                int* line_len, double* norm_speed, double* points, long long* n_points)
{
    PPGrid g; g.nx = nx; g.ny = ny; g.nz = nz; g.off = 0; g.pitch = nx; g.plane = (long long)nx * ny;
    PPParams p;
    p.nsx = density; p.nsy = density / 2; p.nsz = density / 2;
    p.stepx = p.nsx > 1 ? ((double)(nx - 2) - 1.0) / (double)(p.nsx - 1) : 0.0;
    p.stepy = p.nsy > 1 ? ((double)(ny - 2) - 1.0) / (double)(p.nsy - 1) : 0.0;
    p.stepz = p.nsz > 1 ? ((double)(nz - 2) - 1.0) / (double)(p.nsz - 1) : 0.0;
    p.half = integration_steps / 2; p.step_size = step_size; p.change_threshold = threshold;
    for (int k = 0; k < 3; ++k) { p.bmin[k] = bmin[k]; p.bmax[k] = bmax[k]; }
    std::vector<double> slot(3 * (2 * (size_t)p.half + 1));
    int nl = 0; long long np = 0;
    for (int s = 0; s < p.nsx * p.nsy * p.nsz; ++s) {
        int first; double ms;
        const int n = pp_streamline(vx, vy, vz, obs, g, p, s, slot.data(), &first, &ms);
        if (!n) continue;
        line_len[nl] = n;
        const double ns = ms / denom;
        norm_speed[nl] = ns < 1.0 ? ns : 1.0;
        memcpy(points + 3 * np, slot.data() + 3 * first, sizeof(double) * 3 * n);
        ++nl; np += n;
    }
    *n_points = np;
    return nl;
}
