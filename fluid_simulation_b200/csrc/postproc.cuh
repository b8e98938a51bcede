// postproc.cuh -- the arithmetic of the reference's offline viewers as CUDA kernels on the resident fields
// (SURVEY.md 8(f)-4).  The viewers' Qt windows and OpenGL draw calls are not reproduced (no graphics stack on
// this box); what they compute before drawing is:
//   * gui.py:61-79, 265-300   a z-slice colour-mapped to RGB888 (Normalize(clip) -> 256-entry
//                             LinearSegmentedColormap -> uint8) with the obstacle darkened  -> k_pp_slice
//   * GUI/utils.py:40-214     seeded, obstacle-filtered unit-speed streamlines with trilinear sampling in
//                             float64 and their speed colour                              -> k_pp_trace (+ scan / gather)
// One thread per pixel / per seed; the per-seed statements live in postproc_core.h.
#pragma once
#include "handle.cuh"
#include "postproc_core.h"
#include <limits.h>
#include <string.h>
#include <vector>

// ---- the viewers' colour map -----------------------------------------------------------------------
// LinearSegmentedColormap.from_list("density_cmap", [white, lightgreen, green, deepskyblue, blue, darkred,
// red]) (gui.py:38-41, GUI/config.py:14-17) with matplotlib's default N = 256: per channel the anchors sit at
// np.linspace(0, 1, 7) * 255 and entry i interpolates linearly between the two anchors around i
// (matplotlib.colors._create_lookup_table: searchsorted on the left, first / last entry copied).
static void pp_density_cmap(double lut[256][3])
{
    static const unsigned char anchors[7][3] = {
        {0xff, 0xff, 0xff}, {0x90, 0xee, 0x90}, {0x00, 0x80, 0x00}, {0x00, 0xbf, 0xff},
        {0x00, 0x00, 0xff}, {0x8b, 0x00, 0x00}, {0xff, 0x00, 0x00} };
    const int N = 256, K = 7;
    double x[K];
    for (int k = 0; k < K; ++k) {                                  // np.linspace(0, 1, 7) * (N - 1)
        const double v = k == K - 1 ? 1.0 : (double)k * (1.0 / (K - 1)) + 0.0;
        x[k] = v * (N - 1);
    }
    for (int c = 0; c < 3; ++c) {
        double y[K];
        for (int k = 0; k < K; ++k) y[k] = anchors[k][c] / 255.0;
        lut[0][c] = y[0];
        lut[N - 1][c] = y[K - 1];
        for (int i = 1; i < N - 1; ++i) {
            const double xi = (N - 1) * (i == N - 1 ? 1.0 : (double)i * (1.0 / (N - 1)) + 0.0);   // (N-1) * linspace(0,1,N)[i]
            int ind = 0;                                           // np.searchsorted(x, xi), side = "left"
            while (ind < K && x[ind] < xi) ++ind;
            const double dist = (xi - x[ind - 1]) / (x[ind] - x[ind - 1]);
            double v = dist * (y[ind] - y[ind - 1]) + y[ind - 1];
            lut[i][c] = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v);       // np.clip(lut, 0, 1)
        }
    }
}

// ---- z-slice image ---------------------------------------------------------------------------------
// apply_cmap (gui.py:61-69): t = (clip(v, vmin, vmax) - vmin) / (vmax - vmin) in float64, index int(t * 256)
// with t == 1 -> 255, NaN -> the "bad" colour (0, 0, 0); bytes = uint8(lut * 255) were folded on the host.
// overlay_obstacle (gui.py:72-79): where obs > 0.5 the byte becomes uint8(float32(byte) * float32(1 - alpha)).
__global__ void __launch_bounds__(256) k_pp_slice(Grid g, int zl, const float* __restrict__ f, const float* __restrict__ obs,
                                                  double vmin, double vmax, float keep, int overlay,
                                                  const unsigned char* __restrict__ lut, unsigned char* __restrict__ rgb)
{
    const int wp = g.W + 2;
    const int n = wp * g.rows;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int x = i % wp, y = i / wp;
        const long long c = fl_idx(g, x, y, zl);
        const double v = (double)f[c];
        unsigned char r = 0, gg = 0, b = 0;
        if (v == v) {
            double t = v < vmin ? vmin : (v > vmax ? vmax : v);
            t = (t - vmin) / (vmax - vmin);
            t = t * 256.0;
            int k = t == 256.0 ? 255 : (int)t;
            k = k < 0 ? 0 : (k > 255 ? 255 : k);
            r = lut[3 * k]; gg = lut[3 * k + 1]; b = lut[3 * k + 2];
        }
        if (overlay && obs[c] > 0.5f) {
            r = (unsigned char)((float)r * keep); gg = (unsigned char)((float)gg * keep); b = (unsigned char)((float)b * keep);
        }
        rgb[3 * i] = r; rgb[3 * i + 1] = gg; rgb[3 * i + 2] = b;
    }
}

// ---- streamlines -----------------------------------------------------------------------------------
// order-preserving float -> unsigned key, so that atomicMax works on floats of either sign
__device__ __forceinline__ unsigned pp_key(float f)
{
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
static inline float pp_unkey(unsigned k)
{
    const unsigned b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    float f; memcpy(&f, &b, sizeof f); return f;
}

// np.where(obs > 0.5) bounding box (utils.py:127-131) and np.max([vx, vy, vz]) (utils.py:206) over the
// padded volume.  box = { min x, y, z, max x, y, z }; vmax = key of the largest float.
__global__ void __launch_bounds__(256) k_pp_scan_volume(Grid g, const float* __restrict__ vx, const float* __restrict__ vy,
                                                        const float* __restrict__ vz, const float* __restrict__ obs,
                                                        int* __restrict__ box, unsigned* __restrict__ vmax)
{
    const int wp = g.W + 2;
    const long long n = (long long)wp * g.rows * (g.D + 2);
    int lo[3] = { INT_MAX, INT_MAX, INT_MAX }, hi[3] = { -1, -1, -1 };
    float m = -INFINITY;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % wp), y = (int)((i / wp) % g.rows), z = (int)(i / ((long long)wp * g.rows));
        const long long c = fl_idx(g, x, y, z);
        m = fmaxf(m, fmaxf(vx[c], fmaxf(vy[c], vz[c])));
        if (obs[c] > 0.5f) {
            lo[0] = min(lo[0], x); lo[1] = min(lo[1], y); lo[2] = min(lo[2], z);
            hi[0] = max(hi[0], x); hi[1] = max(hi[1], y); hi[2] = max(hi[2], z);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            lo[k] = min(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = max(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(vmax, pp_key(m));
#pragma unroll
        for (int k = 0; k < 3; ++k) { atomicMin(box + k, lo[k]); atomicMax(box + 3 + k, hi[k]); }
    }
}

// one thread per seed; slots = nseeds x (2 * half + 1) x 3 doubles
__global__ void __launch_bounds__(128) k_pp_trace(PPGrid pg, PPParams p, int nseeds, const float* __restrict__ vx,
                                                  const float* __restrict__ vy, const float* __restrict__ vz,
                                                  const float* __restrict__ obs, double* __restrict__ slots,
                                                  int* __restrict__ len, int* __restrict__ first, double* __restrict__ max_speed)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nseeds) return;
    int f0; double ms;
    const int n = pp_streamline(vx, vy, vz, obs, pg, p, s, slots + 3ll * (2 * p.half + 1) * s, &f0, &ms);
    len[s] = n; first[s] = f0; max_speed[s] = ms;
}

// exclusive scan of the accepted lines and of their point counts, in seed order (= the order the
// reference appends them in).  One block; totals[0] = lines, totals[1] = points.
__global__ void __launch_bounds__(1024) k_pp_offsets(int nseeds, const int* __restrict__ len, int* __restrict__ line_id,
                                                     long long* __restrict__ point_off, long long* __restrict__ totals)
{
    __shared__ long long sh_pts[1024];
    __shared__ int sh_lines[1024];
    const int t = threadIdx.x, per = (nseeds + 1023) / 1024;
    const int a = min(nseeds, t * per), b = min(nseeds, a + per);
    long long pts = 0; int lines = 0;
    for (int i = a; i < b; ++i) { pts += len[i]; lines += len[i] > 0; }
    sh_pts[t] = pts; sh_lines[t] = lines;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {                          // Hillis-Steele inclusive scan
        const long long vp = t >= o ? sh_pts[t - o] : 0; const int vl = t >= o ? sh_lines[t - o] : 0;
        __syncthreads();
        sh_pts[t] += vp; sh_lines[t] += vl;
        __syncthreads();
    }
    long long po = sh_pts[t] - pts; int lo = sh_lines[t] - lines;
    for (int i = a; i < b; ++i) { line_id[i] = lo; point_off[i] = po; po += len[i]; lo += len[i] > 0; }
    if (t == 1023) { totals[0] = sh_lines[t]; totals[1] = sh_pts[t]; }
}

// pack the accepted lines: one warp per seed
__global__ void __launch_bounds__(256) k_pp_gather(int nseeds, int slot_pts, const int* __restrict__ len, const int* __restrict__ first,
                                                   const int* __restrict__ line_id, const long long* __restrict__ point_off,
                                                   const double* __restrict__ slots, const double* __restrict__ max_speed,
                                                   double vmax_denominator, int* __restrict__ out_len,
                                                   double* __restrict__ out_norm, double* __restrict__ out_pts)
{
    const int lane = threadIdx.x & 31;
    for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < nseeds; s += (gridDim.x * blockDim.x) >> 5) {
        const int n = len[s];
        if (n == 0) continue;
        const double* src = slots + 3ll * ((long long)slot_pts * s + first[s]);
        double* dst = out_pts + 3 * point_off[s];
        for (int i = lane; i < 3 * n; i += 32) dst[i] = src[i];
        if (lane == 0) {
            out_len[line_id[s]] = n;
            const double ns = max_speed[s] / vmax_denominator;     // utils.py:206
            out_norm[line_id[s]] = ns < 1.0 ? ns : 1.0;            // min(norm_speed, 1.0), :208
        }
    }
}

static PPGrid pp_grid(const Grid& g)
{
    PPGrid p;
    p.nx = g.W + 2; p.ny = g.H + 2; p.nz = g.D + 2;
    p.off = FL_XOFF; p.pitch = g.pitch; p.plane = g.plane;
    return p;
}

static int pp_single_gpu(fluid_handle* h, const char* who)
{
    if (!h) return FLUID_E_ARG;
    if (h->nranks != 1)
        return fl_fail(h, FLUID_E_STATE, "%s: needs the whole volume on one GPU (z-slab runs: view the dumped frames)", who);
    return 0;
}

int fl_render_slice(fluid_handle* h, int field_id, int z, double vmin, double vmax, double obstacle_alpha,
                    unsigned char* rgb_host)
{
    if (int e = pp_single_gpu(h, "fluid_render_slice")) return e;
    const Grid& g = h->g;
    if (field_id < 0 || field_id >= FLUID_NFIELDS || z < 0 || z > g.D + 1 || !rgb_host || !(vmin < vmax))
        return fl_fail(h, FLUID_E_ARG, "fluid_render_slice: bad argument (field %d, z %d, vmin %g, vmax %g)", field_id, z, vmin, vmax);
    double lut[256][3];
    unsigned char lut8[256 * 3];
    pp_density_cmap(lut);
    for (int i = 0; i < 256; ++i)
        for (int c = 0; c < 3; ++c) lut8[3 * i + c] = (unsigned char)(lut[i][c] * 255);   // (rgba[..., :3] * 255).astype(np.uint8)
    const size_t npix = (size_t)(g.W + 2) * g.rows;
    unsigned char *d_lut = nullptr, *d_rgb = nullptr;
    auto release = [&]() { cudaFree(d_lut); cudaFree(d_rgb); };
#define PP_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { release();                       \
        return fl_fail(h, FLUID_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } } while (0)
    PP_CUDA(cudaMalloc(&d_lut, sizeof lut8));
    PP_CUDA(cudaMalloc(&d_rgb, npix * 3));
    PP_CUDA(cudaMemcpyAsync(d_lut, lut8, sizeof lut8, cudaMemcpyHostToDevice, h->stream));
    const float keep = (float)(1.0 - obstacle_alpha);
    k_pp_slice<<<fl_grid_1d((long long)npix), 256, 0, h->stream>>>(g, z, h->field[field_id], h->field[FLUID_OBS], vmin, vmax, keep,
                                                                  obstacle_alpha > 0.0 ? 1 : 0, d_lut, d_rgb);
    PP_CUDA(cudaGetLastError());
    h->launches++;
    PP_CUDA(cudaMemcpyAsync(rgb_host, d_rgb, npix * 3, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(cudaStreamSynchronize(h->stream));
    release();
    return 0;
}

int fl_streamlines(fluid_handle* h, const fluid_streamline_params* sp, int max_lines, long long max_points,
                   int* n_lines, long long* n_points, int* line_len, double* norm_speed, double* points)
{
    if (n_lines) *n_lines = 0;
    if (n_points) *n_points = 0;
    if (int e = pp_single_gpu(h, "fluid_streamlines")) return e;
    if (!sp || !n_lines || !n_points || sp->density < 0 || sp->integration_steps < 0 || max_lines < 0 || max_points < 0)
        return fl_fail(h, FLUID_E_ARG, "fluid_streamlines: bad argument");
    const Grid& g = h->g;
    PPGrid pg = pp_grid(g);
    PPParams p = {};
    p.nsx = sp->density; p.nsy = sp->density / 2; p.nsz = sp->density / 2;
    p.stepx = p.nsx > 1 ? ((double)(pg.nx - 2) - 1.0) / (double)(p.nsx - 1) : 0.0;
    p.stepy = p.nsy > 1 ? ((double)(pg.ny - 2) - 1.0) / (double)(p.nsy - 1) : 0.0;
    p.stepz = p.nsz > 1 ? ((double)(pg.nz - 2) - 1.0) / (double)(p.nsz - 1) : 0.0;
    p.half = sp->integration_steps / 2;
    p.step_size = sp->step_size;
    p.change_threshold = sp->velocity_change_threshold;
    const long long nseeds_ll = (long long)p.nsx * p.nsy * p.nsz;
    const int slot_pts = 2 * p.half + 1;
    if (nseeds_ll > (1 << 24) || nseeds_ll * slot_pts * 24 > (8ll << 30))
        return fl_fail(h, FLUID_E_ARG, "fluid_streamlines: %lld seeds x %d points exceed the scratch limit", nseeds_ll, slot_pts);
    const int nseeds = (int)nseeds_ll;

    int* d_box = nullptr; unsigned* d_vmax = nullptr;
    double *d_slots = nullptr, *d_ms = nullptr, *d_norm = nullptr, *d_pts = nullptr;
    int *d_len = nullptr, *d_first = nullptr, *d_line = nullptr, *d_olen = nullptr;
    long long *d_poff = nullptr, *d_tot = nullptr;
    auto release = [&]() { cudaFree(d_box); cudaFree(d_vmax); cudaFree(d_slots); cudaFree(d_ms); cudaFree(d_norm); cudaFree(d_pts);
                           cudaFree(d_len); cudaFree(d_first); cudaFree(d_line); cudaFree(d_olen); cudaFree(d_poff); cudaFree(d_tot); };
    const int box_init[6] = { INT_MAX, INT_MAX, INT_MAX, -1, -1, -1 };
    PP_CUDA(cudaMalloc(&d_box, sizeof box_init));
    PP_CUDA(cudaMalloc(&d_vmax, sizeof(unsigned)));
    PP_CUDA(cudaMemcpyAsync(d_box, box_init, sizeof box_init, cudaMemcpyHostToDevice, h->stream));
    PP_CUDA(cudaMemsetAsync(d_vmax, 0, sizeof(unsigned), h->stream));
    const long long ncell = (long long)pg.nx * pg.ny * pg.nz;
    k_pp_scan_volume<<<fl_grid_1d(ncell), 256, 0, h->stream>>>(g, h->field[FLUID_VX], h->field[FLUID_VY], h->field[FLUID_VZ],
                                                               h->field[FLUID_OBS], d_box, d_vmax);
    PP_CUDA(cudaGetLastError());
    h->launches++;
    int box[6]; unsigned vkey = 0;
    PP_CUDA(cudaMemcpyAsync(box, d_box, sizeof box, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(cudaMemcpyAsync(&vkey, d_vmax, sizeof vkey, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(cudaStreamSynchronize(h->stream));
    if (box[3] < 0 || nseeds == 0) { release(); return 0; }          // no obstacle: no streamlines (utils.py:134-136)
    const double widen = sp->proximity / 10;                         // utils.py:132-133
    for (int k = 0; k < 3; ++k) { p.bmin[k] = (double)box[k] - widen; p.bmax[k] = (double)box[3 + k] + widen; }
    // np.max([vx, vy, vz]) + 1e-6: a float32 scalar plus a Python float is a float32 sum (NumPy 2 promotion)
    const float vmaxf = pp_unkey(vkey);
    const float denomf = vmaxf + 1e-6f;
    const double denom = (double)denomf;

    PP_CUDA(cudaMalloc(&d_slots, (size_t)nseeds * slot_pts * 3 * sizeof(double)));
    PP_CUDA(cudaMalloc(&d_ms, (size_t)nseeds * sizeof(double)));
    PP_CUDA(cudaMalloc(&d_len, (size_t)nseeds * sizeof(int)));
    PP_CUDA(cudaMalloc(&d_first, (size_t)nseeds * sizeof(int)));
    PP_CUDA(cudaMalloc(&d_line, (size_t)nseeds * sizeof(int)));
    PP_CUDA(cudaMalloc(&d_poff, (size_t)nseeds * sizeof(long long)));
    PP_CUDA(cudaMalloc(&d_tot, 2 * sizeof(long long)));
    k_pp_trace<<<(nseeds + 127) / 128, 128, 0, h->stream>>>(pg, p, nseeds, h->field[FLUID_VX], h->field[FLUID_VY], h->field[FLUID_VZ],
                                                            h->field[FLUID_OBS], d_slots, d_len, d_first, d_ms);
    PP_CUDA(cudaGetLastError());
    k_pp_offsets<<<1, 1024, 0, h->stream>>>(nseeds, d_len, d_line, d_poff, d_tot);
    PP_CUDA(cudaGetLastError());
    h->launches += 2;
    long long tot[2] = { 0, 0 };
    PP_CUDA(cudaMemcpyAsync(tot, d_tot, sizeof tot, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(cudaStreamSynchronize(h->stream));
    *n_lines = (int)tot[0]; *n_points = tot[1];
    if (tot[0] == 0) { release(); return 0; }
    if (tot[0] > max_lines || tot[1] > max_points || !line_len || !norm_speed || !points) {
        release();
        return fl_fail(h, FLUID_E_ARG, "fluid_streamlines: %lld lines / %lld points do not fit the caller's buffers (%d / %lld)",
                       tot[0], tot[1], max_lines, max_points);
    }
    PP_CUDA(cudaMalloc(&d_olen, (size_t)tot[0] * sizeof(int)));
    PP_CUDA(cudaMalloc(&d_norm, (size_t)tot[0] * sizeof(double)));
    PP_CUDA(cudaMalloc(&d_pts, (size_t)tot[1] * 3 * sizeof(double)));
    k_pp_gather<<<fl_grid_1d((long long)nseeds * 32), 256, 0, h->stream>>>(nseeds, slot_pts, d_len, d_first, d_line, d_poff, d_slots,
                                                                          d_ms, denom, d_olen, d_norm, d_pts);
    PP_CUDA(cudaGetLastError());
    h->launches++;
    PP_CUDA(cudaMemcpyAsync(line_len, d_olen, (size_t)tot[0] * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(cudaMemcpyAsync(norm_speed, d_norm, (size_t)tot[0] * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(cudaMemcpyAsync(points, d_pts, (size_t)tot[1] * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(cudaStreamSynchronize(h->stream));
#undef PP_CUDA
    release();
    return 0;
}
