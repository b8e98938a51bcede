// kernels_reduce.cuh -- run()'s statistics (simulation.cpp:73-90) and divergence norms as
// warp-shuffle block reductions; one partial per block, folded by the last block to finish.
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

// one block folds per-block partials (second pass of fluid_stats_static)
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

// ---- one launch: float4 loads over whole planes, the last block to finish folds the partials ----------
// A plane is a contiguous run of pitch*rows floats and a row holds its W+2 cells at float offsets
// FL_XOFF .. FL_XOFF+W+1, so the owned planes are streamed as float4 quads and the pitch padding is masked
// by the quad's position in its row (kept incrementally: no division in the loop).  `ticket` must be zero
// on entry and is zero again on exit.
__device__ __forceinline__ void fl_stat_acc(bool on, float d, float a, float b, float e, double& s, float mn[4], float mx[4])
{
    if (!on) return;
    s += (double)d;
    mn[0] = fminf(mn[0], d); mx[0] = fmaxf(mx[0], d);
    mn[1] = fminf(mn[1], a); mx[1] = fmaxf(mx[1], a);
    mn[2] = fminf(mn[2], b); mx[2] = fmaxf(mx[2], b);
    mn[3] = fminf(mn[3], e); mx[3] = fmaxf(mx[3], e);
}

__global__ void __launch_bounds__(256) k_stats_quads(Grid g, int zl_first, int zl_last,
                                                     const float* __restrict__ dens, const float* __restrict__ vx,
                                                     const float* __restrict__ vy, const float* __restrict__ vz,
                                                     double* partial, unsigned* ticket, double* out)
{
    const int qpr = g.pitch >> 2;                                   // quads per row
    const long long base = (long long)zl_first * g.plane;           // float offset of the first owned plane
    const long long nq = ((long long)(zl_last - zl_first + 1) * g.plane) >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int dcol = (int)(stride % qpr);
    const int last_f = FL_XOFF + g.W + 1;                           // float offset of cell x = W+1 in its row
    const float4* D = reinterpret_cast<const float4*>(dens + base);
    const float4* X = reinterpret_cast<const float4*>(vx + base);
    const float4* Y = reinterpret_cast<const float4*>(vy + base);
    const float4* Z = reinterpret_cast<const float4*>(vz + base);
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int col = (int)(q % qpr);
    double s = 0.0;
    float mn[4] = { FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX }, mx[4] = { -FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (; q < nq; q += stride) {
        const float4 d = D[q], a = X[q], b = Y[q], e = Z[q];
        const int f0 = col << 2;
        fl_stat_acc(f0     >= FL_XOFF && f0     <= last_f, d.x, a.x, b.x, e.x, s, mn, mx);
        fl_stat_acc(f0 + 1 >= FL_XOFF && f0 + 1 <= last_f, d.y, a.y, b.y, e.y, s, mn, mx);
        fl_stat_acc(f0 + 2 >= FL_XOFF && f0 + 2 <= last_f, d.z, a.z, b.z, e.z, s, mn, mx);
        fl_stat_acc(f0 + 3 >= FL_XOFF && f0 + 3 <= last_f, d.w, a.w, b.w, e.w, s, mn, mx);
        col += dcol; if (col >= qpr) col -= qpr;
    }
    block_fold_stats(s, mn, mx, partial + (long long)blockIdx.x * FL_NSTAT);
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        __threadfence();                                            // this block's partials before its ticket
        is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    s = 0.0;
#pragma unroll
    for (int k = 0; k < 4; ++k) { mn[k] = FLT_MAX; mx[k] = -FLT_MAX; }
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) {
        const double* p = partial + (long long)i * FL_NSTAT;     // written by other SMs: read at L2
        s += __ldcg(p);
#pragma unroll
        for (int k = 0; k < 4; ++k) { mn[k] = fminf(mn[k], (float)__ldcg(p + 1 + 2 * k)); mx[k] = fmaxf(mx[k], (float)__ldcg(p + 2 + 2 * k)); }
    }
    __syncthreads();                                                // block_fold_stats reuses its shared array
    block_fold_stats(s, mn, mx, out);
    if (threadIdx.x == 0) *ticket = 0u;
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
