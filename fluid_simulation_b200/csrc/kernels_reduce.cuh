// kernels_reduce.cuh -- run()'s statistics (simulation.cpp:73-90) and divergence norms as
// warp-shuffle block reductions; one partial per block, folded by a single-block second pass.
#pragma once
#include "common.cuh"
#include <float.h>

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_min(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// partial layout per block: [0] sum(dens) [1..2] min/max dens [3..4] v_x [5..6] v_y [7..8] v_z
#define FL_NSTAT 9

__device__ __forceinline__ void block_fold_stats(double s, const float mn[4], const float mx[4],
                                                 double* __restrict__ out)
{
    __shared__ double sh[FL_NSTAT][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    s = warp_sum(s);
    float wmn[4], wmx[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { wmn[k] = warp_min(mn[k]); wmx[k] = warp_max(mx[k]); }
    if (lane == 0) {
        sh[0][warp] = s;
#pragma unroll
        for (int k = 0; k < 4; ++k) { sh[1 + 2 * k][warp] = wmn[k]; sh[2 + 2 * k][warp] = wmx[k]; }
    }
    __syncthreads();
    if (warp == 0) {
        double a = lane < nw ? sh[0][lane] : 0.0;
        a = warp_sum(a);
        float bmn[4], bmx[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            bmn[k] = warp_min(lane < nw ? (float)sh[1 + 2 * k][lane] : FLT_MAX);
            bmx[k] = warp_max(lane < nw ? (float)sh[2 + 2 * k][lane] : -FLT_MAX);
        }
        if (lane == 0) {
            out[0] = a;
#pragma unroll
            for (int k = 0; k < 4; ++k) { out[1 + 2 * k] = bmn[k]; out[2 + 2 * k] = bmx[k]; }
        }
    }
}

// pass 1: every padded cell this slab OWNS (local planes zl_first..zl_last, all x in 0..W+1, y in 0..H+1)
__global__ void __launch_bounds__(256) k_stats_partial(Grid g, int zl_first, int zl_last,
                                                       const float* __restrict__ dens,
                                                       const float* __restrict__ vx,
                                                       const float* __restrict__ vy,
                                                       const float* __restrict__ vz,
                                                       double* __restrict__ partial)
{
    // one x-row of the padded volume per block iteration: the division by the row length happens once
    // per row, not once per cell (the kernel was bound by integer divides)
    const int Wp = g.W + 2;
    const long long nrows = (long long)g.rows * (zl_last - zl_first + 1);
    double s = 0.0;
    float mn[4] = { FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX }, mx[4] = { -FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (long long r = blockIdx.x; r < nrows; r += gridDim.x) {
        const int y = (int)(r % g.rows), zl = zl_first + (int)(r / g.rows);
        const long long c0 = fl_idx(g, 0, y, zl);
        for (int x = threadIdx.x; x < Wp; x += blockDim.x) {
            const long long c = c0 + x;
            float d = dens[c], a = vx[c], b = vy[c], e = vz[c];
            s += (double)d;
            mn[0] = fminf(mn[0], d); mx[0] = fmaxf(mx[0], d);
            mn[1] = fminf(mn[1], a); mx[1] = fmaxf(mx[1], a);
            mn[2] = fminf(mn[2], b); mx[2] = fmaxf(mx[2], b);
            mn[3] = fminf(mn[3], e); mx[3] = fmaxf(mx[3], e);
        }
    }
    block_fold_stats(s, mn, mx, partial + (long long)blockIdx.x * FL_NSTAT);
}

// pass 2: one block folds the per-block partials
__global__ void __launch_bounds__(256) k_stats_final(int nblocks, const double* __restrict__ partial,
                                                     double* __restrict__ out)
{
    double s = 0.0;
    float mn[4] = { FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX }, mx[4] = { -FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (int i = threadIdx.x; i < nblocks; i += blockDim.x) {
        const double* p = partial + (long long)i * FL_NSTAT;
        s += p[0];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            mn[k] = fminf(mn[k], (float)p[1 + 2 * k]);
            mx[k] = fmaxf(mx[k], (float)p[2 + 2 * k]);
        }
    }
    block_fold_stats(s, mn, mx, out);
}

// sum(div^2) and max|div| over interior fluid cells; partial[2*block] = sum, [2*block+1] = max
__global__ void __launch_bounds__(256) k_divnorm_partial(Grid g, const float* __restrict__ div,
                                                         const unsigned char* __restrict__ flags,
                                                         double* __restrict__ partial)
{
    const long long n = (long long)g.W * g.H * g.Dl;
    double s = 0.0; float m = 0.0f;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n;
         t += (long long)gridDim.x * blockDim.x) {
        int x = 1 + (int)(t % g.W);
        long long r = t / g.W;
        int y = 1 + (int)(r % g.H);
        int zl = 1 + (int)(r / g.H);
        long long c = fl_idx(g, x, y, zl);
        if (flags[c] & FL_SOLID) continue;
        float d = div[c];
        s += (double)d * (double)d;
        m = fmaxf(m, fabsf(d));
    }
    __shared__ double shs[32]; __shared__ float shm[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    s = warp_sum(s); m = warp_max(m);
    if (lane == 0) { shs[warp] = s; shm[warp] = m; }
    __syncthreads();
    if (warp == 0) {
        s = warp_sum(lane < nw ? shs[lane] : 0.0);
        m = warp_max(lane < nw ? shm[lane] : 0.0f);
        if (lane == 0) { partial[2 * blockIdx.x] = s; partial[2 * blockIdx.x + 1] = (double)m; }
    }
}
