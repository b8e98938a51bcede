// kernels_basic.cuh -- layout conversion, flag volume, inflow, standalone setBounds, sparse mutators.
#pragma once
#include "common.cuh"

// ---- reference layout <-> pitched HBM layout (host boundary only) ----------------------------
// host volume is the GLOBAL padded array; a slab moves local planes zl <-> global z = zlo-1+zl.
// `staging` holds `nplanes` planes starting at local plane zl_first in reference row order.
__global__ void k_unpack(Grid g, int zl_first, int nplanes, const float* __restrict__ staging,
                         float* __restrict__ field)
{
    const int Wp = g.W + 2;
    long long n = (long long)Wp * g.rows * nplanes;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n;
         t += (long long)gridDim.x * blockDim.x) {
        int x = (int)(t % Wp);
        long long r = t / Wp;
        int y = (int)(r % g.rows);
        int zl = zl_first + (int)(r / g.rows);
        field[fl_idx(g, x, y, zl)] = staging[t];
    }
}

__global__ void k_pack(Grid g, int zl_first, int nplanes, const float* __restrict__ field,
                       float* __restrict__ staging)
{
    const int Wp = g.W + 2;
    long long n = (long long)Wp * g.rows * nplanes;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n;
         t += (long long)gridDim.x * blockDim.x) {
        int x = (int)(t % Wp);
        long long r = t / Wp;
        int y = (int)(r % g.rows);
        int zl = zl_first + (int)(r / g.rows);
        staging[t] = field[fl_idx(g, x, y, zl)];
    }
}

// Frame snapshot (simulation.cpp:140-148): up to five fields repacked to the reference layout in ONE launch.
// One warp per padded x-row (no integer divisions per element): blockDim = (32, 8),
// grid = (ceil(rows * nplanes / 8), fields); reads and writes are both contiguous runs of W + 2 floats, and a lane
// has eight loads in flight before its first store (one load, one store at a time ran at a third of the HBM rate).
struct PackBatch { const float* src[5]; float* dst[5]; };
#define FL_PACK_MLP 8
__global__ void __launch_bounds__(256) k_pack_rows(const __grid_constant__ Grid g, int zl_first, int nplanes,
                                                   const __grid_constant__ PackBatch pb)
{
    const int r = blockIdx.x * blockDim.y + threadIdx.y;              // padded row index inside the packed planes
    if (r >= g.rows * nplanes) return;
    const int zl = zl_first + r / g.rows, y = r - (r / g.rows) * g.rows;
    const int Wp = g.W + 2;
    const float* __restrict__ s = pb.src[blockIdx.y] + fl_idx(g, 0, y, zl);
    float* __restrict__ d = pb.dst[blockIdx.y] + (long long)r * Wp;
    for (int x0 = threadIdx.x; x0 < Wp; x0 += 32 * FL_PACK_MLP) {
        float v[FL_PACK_MLP];
#pragma unroll
        for (int k = 0; k < FL_PACK_MLP; ++k) v[k] = (x0 + 32 * k < Wp) ? __ldg(s + x0 + 32 * k) : 0.0f;
#pragma unroll
        for (int k = 0; k < FL_PACK_MLP; ++k) if (x0 + 32 * k < Wp) d[x0 + 32 * k] = v[k];
    }
}

// ---- flag volume (once per obstacle upload) ---------------------------------------------------
// Encodes exactly the obs tests the reference performs: `== 1.0f` for solid, `== 0.0f` plus the
// 1..dim range test for "open" neighbours (simulation.cpp:222, 232-238, 307-312, 329-355).
// obs has valid halo planes in slab mode, so the +-z tests see the neighbouring slab's mask.
// obs / flags are the ALLOCATION bases (plane -FL_ZPAD first); index t runs over the whole allocation.
__global__ void k_build_flags(Grid g, const float* __restrict__ obs, unsigned char* __restrict__ flags)
{
    long long n = g.cells;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n;
         t += (long long)gridDim.x * blockDim.x) {
        long long r = t;
        int zl = (int)(r / g.plane); r -= (long long)zl * g.plane; zl -= FL_ZPAD;
        int y = (int)(r / g.pitch);
        int x = (int)(r - (long long)y * g.pitch) - FL_XOFF;
        int gz = g.zlo + zl - 1;
        unsigned f = 0;
        // owned planes and, on a z-slab, the halo planes the temporally blocked sweep zeroes cells on (their own
        // z-neighbours are held too); everything outside the tunnel's interior stays 0
        if (x >= 1 && x <= g.W && y >= 1 && y <= g.H && gz >= 1 && gz <= g.D && zl >= 2 - FL_HALO && zl <= g.Dl + FL_HALO - 1) {
            const float o = obs[t];
            const bool inxp = x + 1 <= g.W, inxm = x - 1 >= 1, inyp = y + 1 <= g.H, inym = y - 1 >= 1;
            const bool inzp = gz + 1 <= g.D, inzm = gz - 1 >= 1;
            const float oxp = obs[t + 1], oxm = obs[t - 1];
            const float oyp = obs[t + g.pitch], oym = obs[t - g.pitch];
            const float ozp = obs[t + g.plane], ozm = obs[t - g.plane];
            if (o == 1.0f) f |= FL_SOLID;
            else if ((inxp && oxp == 1.0f) || (inxm && oxm == 1.0f) || (inyp && oyp == 1.0f) ||
                     (inym && oym == 1.0f) || (inzp && ozp == 1.0f) || (inzm && ozm == 1.0f))
                f |= FL_NOSLIP;
            if (inxp && oxp == 0.0f) f |= FL_OPEN_XP;
            if (inxm && oxm == 0.0f) f |= FL_OPEN_XM;
            if (inyp && oyp == 0.0f) f |= FL_OPEN_YP;
            if (inym && oym == 0.0f) f |= FL_OPEN_YM;
            if (inzp && ozp == 0.0f) f |= FL_OPEN_ZP;
            if (inzm && ozm == 0.0f) f |= FL_OPEN_ZM;
        }
        flags[t] = (unsigned char)f;
    }
}

// ---- inlet on the x = 1 face (no setBounds follows in the reference) --------------------------
// run(): dens(1,j,k) += 0.001f            simulation.cpp:65-67
__global__ void k_inflow_density(Grid g, float* __restrict__ dens, float amount)
{
    int y = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    int zl = (int)blockIdx.y - (FL_HALO - 1);    // 1-FL_HALO .. Dl+FL_HALO: a slab also keeps its halo planes current
    int gz = g.zlo + zl - 1;
    if (y <= g.H && gz >= 1 && gz <= g.D) dens[fl_idx(g, 1, y, zl)] += amount;
}

// step(): v(1,j,k) = (speed, 0, 0)        simulation.cpp:103-105
__global__ void k_inflow_velocity(Grid g, float* __restrict__ vx, float* __restrict__ vy,
                                  float* __restrict__ vz, float speed)
{
    int y = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    int zl = (int)blockIdx.y - (FL_HALO - 1);
    int gz = g.zlo + zl - 1;
    if (y <= g.H && gz >= 1 && gz <= g.D) {
        long long c = fl_idx(g, 1, y, zl);
        vx[c] = speed; vy[c] = 0.0f; vz[c] = 0.0f;
    }
}

// ---- standalone setBounds(b, field), simulation.cpp:183-246 -----------------------------------
// One thread owns each interior quad: it reads the stored (pre-zero) value once, derives the
// ghosts from that register copy, then writes the masked value -- no inter-thread ordering needed.
__global__ void __launch_bounds__(256) k_set_bounds(Grid g, float* __restrict__ field,
                                                    const unsigned char* __restrict__ flags, int b)
{
    QuadPos p = fl_quad_pos(g, 1 + blockIdx.z);
    if (!p.valid) return;
    float4 c = ld4(field + fl_idx(g, p.x0, p.y, p.zl));
    float raw[4] = { c.x, c.y, c.z, c.w };
    fl_store_bounded(g, field, flags, b, p.x0, p.y, p.zl, raw);
}

// ---- edge / corner ghosts --------------------------------------------------------------------
// setBounds never writes cells with two or more ghost coordinates (simulation.cpp:187-215), so an
// in-place solver keeps whatever they held.  Ping-pong (Jacobi) partners get them copied once per
// solve so the result is independent of which buffer the last sweep landed in.
__global__ void k_copy_edges(Grid g, const float* __restrict__ src, float* __restrict__ dst)
{
    const int per = 2 * (g.W + 2) + 2 * g.H;                 // perimeter cells of one plane
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int zl = blockIdx.y;
    if (t >= per) return;
    int x, y;
    if (t < g.W + 2)            { x = t; y = 0; }
    else if (t < 2 * (g.W + 2)) { x = t - (g.W + 2); y = g.H + 1; }
    else if (t < 2 * (g.W + 2) + g.H) { x = 0; y = 1 + t - 2 * (g.W + 2); }
    else                        { x = g.W + 1; y = 1 + t - 2 * (g.W + 2) - g.H; }
    const int gz = g.zlo + zl - 1;
    const int nghost = (x == 0 || x == g.W + 1) + (y == 0 || y == g.H + 1) + (gz == 0 || gz == g.D + 1);
    if (nghost >= 2) { long long c = fl_idx(g, x, y, zl); dst[c] = src[c]; }
}

// ---- sparse mutators (addObstacle / addDensity / setVelocity, simulation.cpp:155-178) --------
// xyz are GLOBAL 0..dim+1 coordinates; cells outside this slab's planes (incl. halos) are skipped.
__global__ void k_add_density(Grid g, float* __restrict__ dens, int n, const int* __restrict__ xyz,
                              const float* __restrict__ amount)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    int zl = xyz[3 * t + 2] - g.zlo + 1;
    if (zl < g.zl_lo || zl > g.zl_hi) return;
    atomicAdd(&dens[fl_idx(g, xyz[3 * t], xyz[3 * t + 1], zl)], amount[t]);
}

__global__ void k_set_velocity(Grid g, float* __restrict__ vx, float* __restrict__ vy,
                               float* __restrict__ vz, int n, const int* __restrict__ xyz,
                               const float* __restrict__ v)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    int zl = xyz[3 * t + 2] - g.zlo + 1;
    if (zl < g.zl_lo || zl > g.zl_hi) return;
    long long c = fl_idx(g, xyz[3 * t], xyz[3 * t + 1], zl);
    vx[c] = v[3 * t]; vy[c] = v[3 * t + 1]; vz[c] = v[3 * t + 2];
}

__global__ void k_add_obstacles(Grid g, float* __restrict__ obs, int n, const int* __restrict__ xyz)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    int zl = xyz[3 * t + 2] - g.zlo + 1;
    if (zl < g.zl_lo || zl > g.zl_hi) return;
    obs[fl_idx(g, xyz[3 * t], xyz[3 * t + 1], zl)] = 1.0f;
}
