/* oracle/postproc_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the arithmetic of the reference's offline viewers, the checker
 * for csrc/postproc.cuh:
 *   - GUI/utils.py:40-81   _interpolate_scalar / _interpolate_vector
 *   - GUI/utils.py:83-115  _integrate_streamline_part
 *   - GUI/utils.py:118-214 generate_streamlines (seeding, culling, filters, speed colour argument)
 *   - gui.py:61-79         apply_cmap + overlay_obstacle (the colour table itself is restated in
 *                          oracle/postproc.py from matplotlib's published algorithm)
 * It follows the Python statement by statement: lists are built per direction, reversed and concatenated,
 * and the filters run over the concatenated lists afterwards.  float64 throughout on float32 volumes, as
 * NumPy does for `np.float32 * np.float64`; no contraction (-ffp-contract=off).
 *
 * Pinning: tests/golden/make_postproc_golden.py runs the reference's own GUI/utils.py (imported from
 * /root/reference with `skimage` and `config` stubbed: neither is installed, config needs matplotlib) on
 * seeded fields and commits its output; tests/test_postproc.py compares this file against it.  The only
 * arithmetic that can differ is np.linalg.norm (BLAS ddot of three products, summation order and FMA use
 * unspecified), hence a 1e-9 tolerance there instead of bit equality.  The colour-mapped slice has no
 * reference run behind it (matplotlib is absent): PARITY UNPINNED for apply_cmap.
 *
 * Volumes are in the frame-file layout: index x + y*nx + z*nx*ny over the padded grid (nx = W+2 ...);
 * GUI/main_window.py:204,227-230 transposes them to [x, y, z], which is what the coordinates below mean. */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct { const float* a; int nx, ny, nz; } vol_t;

static double at(const vol_t* g, int x, int y, int z)
{
    return (double)g->a[(size_t)x + (size_t)y * g->nx + (size_t)z * g->nx * g->ny];
}

static double clip(double v, double lo, double hi) { return fmin(fmax(v, lo), hi); }

/* utils.py:40-73 */
static double interpolate_scalar(const vol_t* g, double x, double y, double z)
{
    x = clip(x, 0, g->nx - 1.001);
    y = clip(y, 0, g->ny - 1.001);
    z = clip(z, 0, g->nz - 1.001);
    int x0 = (int)x, y0 = (int)y, z0 = (int)z;
    int x1 = x0 + 1, y1 = y0 + 1, z1 = z0 + 1;
    double xd = x - x0, yd = y - y0, zd = z - z0;
    double c000 = at(g, x0, y0, z0), c100 = at(g, x1, y0, z0), c010 = at(g, x0, y1, z0), c110 = at(g, x1, y1, z0);
    double c001 = at(g, x0, y0, z1), c101 = at(g, x1, y0, z1), c011 = at(g, x0, y1, z1), c111 = at(g, x1, y1, z1);
    double c00 = c000 * (1 - xd) + c100 * xd;
    double c01 = c001 * (1 - xd) + c101 * xd;
    double c10 = c010 * (1 - xd) + c110 * xd;
    double c11 = c011 * (1 - xd) + c111 * xd;
    double c0 = c00 * (1 - yd) + c10 * yd;
    double c1 = c01 * (1 - yd) + c11 * yd;
    return c0 * (1 - zd) + c1 * zd;
}

/* utils.py:75-81 */
static void interpolate_vector(const vol_t* vx, const vol_t* vy, const vol_t* vz, const double p[3], double out[3])
{
    out[0] = interpolate_scalar(vx, p[0], p[1], p[2]);
    out[1] = interpolate_scalar(vy, p[0], p[1], p[2]);
    out[2] = interpolate_scalar(vz, p[0], p[1], p[2]);
}

static double norm3(const double v[3]) { return sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

/* utils.py:83-115; pts / vel hold max_steps + 1 triples; returns the list length */
static int integrate_part(const double start[3], const vol_t* vx, const vol_t* vy, const vol_t* vz, const vol_t* obs,
                          int max_steps, double direction, double step_size, double* pts, double* vel)
{
    int n = 1;
    double pos[3] = { start[0], start[1], start[2] };
    memcpy(pts, start, sizeof pos);
    interpolate_vector(vx, vy, vz, start, vel);
    for (int it = 0; it < max_steps; ++it) {
        double vec[3];
        interpolate_vector(vx, vy, vz, pos, vec);
        double speed = norm3(vec);
        if (speed < 1e-6) break;
        for (int k = 0; k < 3; ++k) {
            double step = direction * (vec[k] / speed) * step_size;
            pos[k] = pos[k] + step;
        }
        if (isnan(pos[0]) || isnan(pos[1]) || isnan(pos[2]) || isinf(pos[0]) || isinf(pos[1]) || isinf(pos[2])) break;
        if (!(1 <= pos[0] && pos[0] < vx->nx - 1 && 1 <= pos[1] && pos[1] < vx->ny - 1 && 1 <= pos[2] && pos[2] < vx->nz - 1))
            break;
        if (interpolate_scalar(obs, pos[0], pos[1], pos[2]) > 0.5) break;
        memcpy(pts + 3 * n, pos, sizeof pos);
        memcpy(vel + 3 * n, vec, sizeof vec);
        ++n;
    }
    return n;
}

/* np.linspace(start, stop, num): arange(num) * ((stop - start) / (num - 1)) + start, last element = stop */
static void linspace(double start, double stop, int num, double* out)
{
    if (num <= 0) return;
    if (num == 1) { out[0] = start; return; }
    double step = (stop - start) / (num - 1);
    for (int i = 0; i < num; ++i) out[i] = (double)i * step + start;
    out[num - 1] = stop;
}

/* utils.py:118-214.  Returns the number of accepted lines, or -1 if the caller's buffers are too small;
 * *n_points = total points.  line_len / norm_speed: one entry per line; points: packed x, y, z.
 * norm_speed is min(max_speed / (np.max([vx, vy, vz]) + 1e-6), 1.0), the colour-map argument (:206-208). */
int oracle_streamlines(const float* vx_a, const float* vy_a, const float* vz_a, const float* obs_a, int nx, int ny, int nz,
                       int density, double proximity, int integration_steps, double step_size, double change_threshold,
                       int max_lines, long long max_points, int* line_len, double* norm_speed, double* points,
                       long long* n_points)
{
    vol_t vx = { vx_a, nx, ny, nz }, vy = { vy_a, nx, ny, nz }, vz = { vz_a, nx, ny, nz }, obs = { obs_a, nx, ny, nz };
    *n_points = 0;
    /* :127-136 bounding box of obs > 0.5 */
    int lo[3] = { nx, ny, nz }, hi[3] = { -1, -1, -1 };
    float vmax = -INFINITY;
    for (int z = 0; z < nz; ++z) for (int y = 0; y < ny; ++y) for (int x = 0; x < nx; ++x) {
        size_t i = (size_t)x + (size_t)y * nx + (size_t)z * nx * ny;
        if (vx_a[i] > vmax) vmax = vx_a[i];
        if (vy_a[i] > vmax) vmax = vy_a[i];
        if (vz_a[i] > vmax) vmax = vz_a[i];
        if (obs_a[i] > 0.5f) {
            const int c[3] = { x, y, z };
            for (int k = 0; k < 3; ++k) { if (c[k] < lo[k]) lo[k] = c[k]; if (c[k] > hi[k]) hi[k] = c[k]; }
        }
    }
    if (hi[0] < 0) return 0;
    double omin[3], omax[3];
    for (int k = 0; k < 3; ++k) { omin[k] = lo[k] - proximity / 10; omax[k] = hi[k] + proximity / 10; }
    /* np.max([vx, vy, vz]) is a float32; adding the Python float 1e-6 is a float32 sum under NumPy 2 */
    const float denom_f = vmax + 1e-6f;

    const int nsx = density, nsy = density / 2, nsz = density / 2, half = integration_steps / 2;
    double* xs = malloc(sizeof(double) * (nsx + 1)); double* ys = malloc(sizeof(double) * (nsy + 1)); double* zs = malloc(sizeof(double) * (nsz + 1));
    linspace(1, nx - 2, nsx, xs); linspace(1, ny - 2, nsy, ys); linspace(1, nz - 2, nsz, zs);
    const size_t cap = (size_t)half + 1;
    double* bp = malloc(cap * 24); double* bv = malloc(cap * 24); double* fp = malloc(cap * 24); double* fv = malloc(cap * 24);
    double* full_p = malloc(2 * cap * 24); double* full_v = malloc(2 * cap * 24);
    int nlines = 0; long long npts = 0; int overflow = 0;

    for (int iz = 0; iz < nsz; ++iz) for (int iy = 0; iy < nsy; ++iy) for (int ix = 0; ix < nsx; ++ix) {
        const double start[3] = { xs[ix], ys[iy], zs[iz] };
        if (start[0] < omin[0] || start[0] > omax[0] || start[1] < omin[1] || start[1] > omax[1] ||
            start[2] < omin[2] || start[2] > omax[2]) continue;                                   /* :150-153 */
        if (at(&obs, (int)start[0], (int)start[1], (int)start[2]) > 0.5) continue;               /* :156-157 */
        int nb = integrate_part(start, &vx, &vy, &vz, &obs, half, -1.0, step_size, bp, bv);
        int nf = integrate_part(start, &vx, &vy, &vz, &obs, half, 1.0, step_size, fp, fv);
        /* backward[::-1][:-1] + forward                                                            :169-170 */
        int n = 0;
        for (int i = nb - 1; i >= 1; --i, ++n) { memcpy(full_p + 3 * n, bp + 3 * i, 24); memcpy(full_v + 3 * n, bv + 3 * i, 24); }
        for (int i = 0; i < nf; ++i, ++n) { memcpy(full_p + 3 * n, fp + 3 * i, 24); memcpy(full_v + 3 * n, fv + 3 * i, 24); }
        if (n <= 5) continue;                                                                     /* :173-174 */
        double max_change = 0.0;                                                                  /* :177-183 */
        for (int i = 1; i < n; ++i) {
            double d[3] = { full_v[3 * i] - full_v[3 * i - 3], full_v[3 * i + 1] - full_v[3 * i - 2], full_v[3 * i + 2] - full_v[3 * i - 1] };
            double change = norm3(d);
            if (change > max_change) max_change = change;
        }
        if (max_change < change_threshold) continue;
        int near = 0;                                                                             /* :186-198 */
        for (int i = 0; i < n; i += 3) {
            const double* pt = full_p + 3 * i;
            if (omin[0] <= pt[0] && pt[0] <= omax[0] && omin[1] <= pt[1] && pt[1] <= omax[1] && omin[2] <= pt[2] && pt[2] <= omax[2]) {
                near = 1; break;
            }
        }
        if (!near) continue;
        double max_speed = 0.0;                                                                   /* :203-204 */
        for (int i = 0; i < n; ++i) { double s = norm3(full_v + 3 * i); if (i == 0 || s > max_speed) max_speed = s; }
        double ns = max_speed / (double)denom_f;                                                  /* :206 */
        if (1.0 < ns) ns = 1.0;                                                                   /* min(norm_speed, 1.0) */
        if (nlines < max_lines && npts + n <= max_points) {
            line_len[nlines] = n; norm_speed[nlines] = ns;
            memcpy(points + 3 * npts, full_p, (size_t)n * 24);
        } else overflow = 1;
        ++nlines; npts += n;
    }
    free(xs); free(ys); free(zs); free(bp); free(bv); free(fp); free(fv); free(full_p); free(full_v);
    *n_points = npts;
    return overflow ? -1 : nlines;
}

/* gui.py:61-79 on one z plane: f / obs point at the plane (ny rows of nx floats); lut = 256 x 3 float64 */
void oracle_render_slice(const float* f, const float* obs, int nx, int ny, double vmin, double vmax, double alpha,
                         const double* lut, unsigned char* rgb)
{
    for (int i = 0; i < nx * ny; ++i) {
        unsigned char px[3] = { 0, 0, 0 };                 /* NaN -> the colour map's "bad" colour (0, 0, 0, 0) */
        double v = (double)f[i];
        if (!isnan(v)) {
            double t = (clip(v, vmin, vmax) - vmin) / (vmax - vmin);     /* Normalize(vmin, vmax, clip=True) */
            double xa = t * 256;                                         /* Colormap.__call__: xa *= N        */
            int k = xa == 256 ? 255 : (int)xa;                           /* xa[xa == N] = N - 1; astype(int)  */
            for (int c = 0; c < 3; ++c) px[c] = (unsigned char)(lut[3 * k + c] * 255);   /* (rgba * 255).astype(uint8) */
        }
        if (obs[i] > 0.5f)                                               /* overlay_obstacle, float32 product */
            for (int c = 0; c < 3; ++c) px[c] = (unsigned char)((float)px[c] * (float)(1 - alpha));
        memcpy(rgb + 3 * i, px, 3);
    }
}
