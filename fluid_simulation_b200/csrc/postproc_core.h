// postproc_core.h -- per-seed streamline arithmetic of the reference's 3-D viewer
// (/root/reference/GUI/utils.py:40-115 trilinear sampling + unit-speed integration, :118-214 seeding and
// filters), written once for the CUDA kernel in postproc.cuh.  It is plain C++ without any CUDA construct
// so that tests/host_emul can also compile it with g++ and run the very same statements on the host as a
// check of this source where no GPU is at hand (not a product path: the library only calls it from kernels).
//
// The reference computes in float64 on float32 volumes indexed [x, y, z] over the PADDED grid
// (GUI/main_window.py:204, 227-230 transpose the (z, y, x) frames); so does this code.  No FMA: the
// library is built with --fmad=false and the emulation with -ffp-contract=off.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define PP_HD __host__ __device__ __forceinline__
#else
#define PP_HD static inline
#endif

// a padded float volume addressed as f[off + x + y*pitch + z*plane]; nx, ny, nz = W+2, H+2, D+2
struct PPGrid {
    int nx, ny, nz;
    long long off, pitch, plane;
};

// parameters of generate_streamlines (GUI/config.py:19-24) and what the host derives from them
struct PPParams {
    int nsx, nsy, nsz;            // seeds per axis: DENSITY, DENSITY // 2, DENSITY // 2   (utils.py:139-141)
    double stepx, stepy, stepz;   // np.linspace step (stop - start) / (num - 1); 0 when num == 1
    int half;                     // max_length // 2 integration steps per direction        (utils.py:160-166)
    double step_size;             // INTEGRATION_STEP_SIZE
    double change_threshold;      // VELOCITY_CHANGE_THRESHOLD
    double bmin[3], bmax[3];      // obstacle bounding box widened by PROXIMITY / 10          (utils.py:127-133)
};

PP_HD double pp_at(const float* f, const PPGrid& g, int x, int y, int z)
{
    return (double)f[g.off + x + (long long)y * g.pitch + (long long)z * g.plane];
}

// np.clip(v, 0, hi) for one float64
PP_HD double pp_clip(double v, double hi)
{
    v = v < 0.0 ? 0.0 : v;
    return v > hi ? hi : v;
}

// _interpolate_scalar, utils.py:40-73: x-then-y-then-z lerps, each `c0 * (1 - t) + c1 * t`
PP_HD double pp_interp(const float* f, const PPGrid& g, double x, double y, double z)
{
    x = pp_clip(x, (double)g.nx - 1.001);
    y = pp_clip(y, (double)g.ny - 1.001);
    z = pp_clip(z, (double)g.nz - 1.001);
    const int x0 = (int)x, y0 = (int)y, z0 = (int)z;
    const int x1 = x0 + 1, y1 = y0 + 1, z1 = z0 + 1;
    const double xd = x - x0, yd = y - y0, zd = z - z0;
    const double c000 = pp_at(f, g, x0, y0, z0), c100 = pp_at(f, g, x1, y0, z0);
    const double c010 = pp_at(f, g, x0, y1, z0), c110 = pp_at(f, g, x1, y1, z0);
    const double c001 = pp_at(f, g, x0, y0, z1), c101 = pp_at(f, g, x1, y0, z1);
    const double c011 = pp_at(f, g, x0, y1, z1), c111 = pp_at(f, g, x1, y1, z1);
    const double c00 = c000 * (1 - xd) + c100 * xd;
    const double c01 = c001 * (1 - xd) + c101 * xd;
    const double c10 = c010 * (1 - xd) + c110 * xd;
    const double c11 = c011 * (1 - xd) + c111 * xd;
    const double c0 = c00 * (1 - yd) + c10 * yd;
    const double c1 = c01 * (1 - yd) + c11 * yd;
    return c0 * (1 - zd) + c1 * zd;
}

// np.linalg.norm of a 3-vector: sqrt(x.x), products summed left to right
PP_HD double pp_norm3(double a, double b, double c) { return sqrt(a * a + b * b + c * c); }

// running filters over the velocities a streamline collects (utils.py:176-183, 203-204)
struct PPLineStat { double max_change, max_speed; };

// _integrate_streamline_part, utils.py:83-115.  pts receives the points at pts[3 * (k * dir_stride)],
// k = 0 .. n-1 (k = 0 is the seed); dir_stride = -1 lays the backward half out in reverse, as
// `backward_points[::-1]` does.  Returns n.
PP_HD int pp_trace_part(const float* vx, const float* vy, const float* vz, const float* obs, const PPGrid& g,
                        double sx, double sy, double sz, int max_steps, double direction, double step_size,
                        double* pts, int dir_stride, PPLineStat& st)
{
    double p0 = sx, p1 = sy, p2 = sz;
    pts[0] = p0; pts[1] = p1; pts[2] = p2;
    // streamline_velocities[0] = velocity at the seed (:86)
    double l0 = pp_interp(vx, g, p0, p1, p2), l1 = pp_interp(vy, g, p0, p1, p2), l2 = pp_interp(vz, g, p0, p1, p2);
    {
        const double s = pp_norm3(l0, l1, l2);
        if (s > st.max_speed) st.max_speed = s;
    }
    int n = 1;
    for (int it = 0; it < max_steps; ++it) {
        const double u = pp_interp(vx, g, p0, p1, p2), v = pp_interp(vy, g, p0, p1, p2), w = pp_interp(vz, g, p0, p1, p2);
        const double speed = pp_norm3(u, v, w);
        if (speed < 1e-6) break;                                                        // :92-93
        // step = direction * (vec / speed) * STEP; pos = pos + step                       :96-97
        const double q0 = p0 + direction * (u / speed) * step_size;
        const double q1 = p1 + direction * (v / speed) * step_size;
        const double q2 = p2 + direction * (w / speed) * step_size;
        p0 = q0; p1 = q1; p2 = q2;
        if (!isfinite(p0) || !isfinite(p1) || !isfinite(p2)) break;                     // :100-101
        if (!(1.0 <= p0 && p0 < (double)(g.nx - 1) && 1.0 <= p1 && p1 < (double)(g.ny - 1) &&
              1.0 <= p2 && p2 < (double)(g.nz - 1))) break;                             // :104-107
        if (pp_interp(obs, g, p0, p1, p2) > 0.5) break;                                 // :110-111
        double* o = pts + 3 * (long long)(n * dir_stride);
        o[0] = p0; o[1] = p1; o[2] = p2;                                                // :113
        // :114 appends `vec`, the velocity the step was taken with
        const double c = pp_norm3(u - l0, v - l1, w - l2);
        if (c > st.max_change) st.max_change = c;
        if (speed > st.max_speed) st.max_speed = speed;
        l0 = u; l1 = v; l2 = w;
        ++n;
    }
    return n;
}

// np.linspace(1, stop, num)[i] with step = (stop - 1) / (num - 1): i * step + 1, last element = stop
PP_HD double pp_seed(int i, int num, double step, double stop)
{
    if (num > 1 && i == num - 1) return stop;
    return (double)i * step + 1.0;
}

PP_HD bool pp_in_box(const PPParams& p, double x, double y, double z)
{
    return p.bmin[0] <= x && x <= p.bmax[0] && p.bmin[1] <= y && y <= p.bmax[1] && p.bmin[2] <= z && z <= p.bmax[2];
}

// One seed of generate_streamlines' triple loop (utils.py:144-211).  slot holds 2 * half + 1 points; the
// seed goes to slot[half], the backward half before it in reverse, the forward half after it, which is
// `backward_points[::-1][:-1] + forward_points`.  Returns the number of points of the accepted line
// (0 = culled or filtered out); *first = index of its first point in the slot, *max_speed = max |v| on it.
PP_HD int pp_streamline(const float* vx, const float* vy, const float* vz, const float* obs, const PPGrid& g,
                        const PPParams& p, int seed, double* slot, int* first, double* max_speed)
{
    const int ix = seed % p.nsx, iy = (seed / p.nsx) % p.nsy, iz = seed / (p.nsx * p.nsy);   // z outer, x inner (:144-146)
    const double sx = pp_seed(ix, p.nsx, p.stepx, (double)(g.nx - 2));
    const double sy = pp_seed(iy, p.nsy, p.stepy, (double)(g.ny - 2));
    const double sz = pp_seed(iz, p.nsz, p.stepz, (double)(g.nz - 2));
    *first = p.half; *max_speed = 0.0;
    if (sx < p.bmin[0] || sx > p.bmax[0] || sy < p.bmin[1] || sy > p.bmax[1] || sz < p.bmin[2] || sz > p.bmax[2])
        return 0;                                                                        // :150-153
    if (pp_at(obs, g, (int)sx, (int)sy, (int)sz) > 0.5) return 0;                        // :156-157
    PPLineStat st; st.max_change = 0.0; st.max_speed = 0.0;
    double* mid = slot + 3 * (long long)p.half;
    const int nb = pp_trace_part(vx, vy, vz, obs, g, sx, sy, sz, p.half, -1.0, p.step_size, mid, -1, st);
    const int nf = pp_trace_part(vx, vy, vz, obs, g, sx, sy, sz, p.half, 1.0, p.step_size, mid, 1, st);
    const int len = nb - 1 + nf;
    const int f0 = p.half - (nb - 1);
    if (len <= 5) return 0;                                                              // :173-174
    if (st.max_change < p.change_threshold) return 0;                                    // :177-183
    bool near = false;                                                                   // :186-195, every 3rd point
    for (int i = 0; i < len && !near; i += 3) {
        const double* q = slot + 3 * (long long)(f0 + i);
        near = pp_in_box(p, q[0], q[1], q[2]);
    }
    if (!near) return 0;
    *first = f0; *max_speed = st.max_speed;
    return len;
}
