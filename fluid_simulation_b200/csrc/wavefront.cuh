// wavefront.cuh -- FLUID_SOLVER_LEX: the reference's in-place lexicographic Gauss-Seidel
// (simulation.cpp:258-272, x outer / y / z inner) executed as a pipelined hyperplane wavefront.
//
// In the 1-thread reference, sweep k at cell (i,j,l) reads
//   * "lower" neighbours (i-1,j,l) (i,j-1,l) (i,j,l-1): already updated in THIS sweep (raw values,
//     solids not yet zeroed -- setBounds only runs after the sweep), and
//   * "upper" neighbours (i+1,..): values of sweep k-1 AFTER its setBounds (masked, ghosts set).
// All cells on a hyperplane s = i+j+l are therefore independent, (k,s) depends on (k,s-1),
// (k-1,s+1) and (k-1,s), and (k,s) can run at time t = s + 2k: one launch per t updates up to
// `acc` hyperplanes of different sweeps at once.  Storage holds the RAW value of the latest sweep
// of each cell; the zero mask of sweep k-1's setBounds is applied when an upper neighbour is
// read (not for k = 0, which reads the caller's field as is); each face ghost is only ever read
// by its single adjacent interior cell, whose thread rewrites it right after updating that cell.
// Writes at time t go to planes s, reads come from planes s+-1 and distinct sweeps are an even
// number of planes apart, so there are no intra-launch hazards.  A final k_set_bounds applies the
// last mask.  The float expression is the same fl_relax, so the result is bit-identical to the
// reference at OMP_NUM_THREADS=1 (tests/test_gpu_lex.py).
#pragma once
#include "handle.cuh"
#include "kernels_basic.cuh"
#include "kernels_stencil.cuh"

struct LexBatch { float* x[3]; const float* rhs[3]; int b[3]; };

__global__ void __launch_bounds__(256) k_lex_wavefront(Grid g, LexBatch lb,
                                                       const unsigned char* __restrict__ flags,
                                                       float a, float c_recip, int t, int kmin)
{
    const int k = kmin + blockIdx.y;
    const int s = t - 2 * k;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= g.H * g.D) return;
    const int j = 1 + tid % g.H, l = 1 + tid / g.H;
    const int i = s - j - l;
    if (i < 1 || i > g.W) return;
    const int batch = blockIdx.z;
    float* __restrict__ x = lb.x[batch];
    const int b = lb.b[batch];
    const unsigned zm = fl_zero_mask(b);
    const long long c = fl_idx(g, i, j, l);

    // upper neighbours: post-setBounds values of the previous sweep
    float xp = x[c + 1], yp = x[c + g.pitch], zp = x[c + g.plane];
    if (k > 0) {
        if (i + 1 <= g.W && (flags[c + 1] & zm)) xp = 0.0f;
        if (j + 1 <= g.H && (flags[c + g.pitch] & zm)) yp = 0.0f;
        if (l + 1 <= g.D && (flags[c + g.plane] & zm)) zp = 0.0f;
    }
    // lower neighbours: raw values of this sweep (or stored ghosts)
    const float xm = x[c - 1], ym = x[c - g.pitch], zmv = x[c - g.plane];
    const float raw = fl_relax(lb.rhs[batch][c], a, c_recip, xp, xm, yp, ym, zp, zmv);
    x[c] = raw;
    if (i == 1)   x[c - 1] = (b == 1) ? -raw : raw;          // simulation.cpp:189
    if (i == g.W) x[c + 1] = raw;                            // :191 outflow copy
    if (j == 1)   x[c - g.pitch] = (b == 2) ? -raw : raw;    // :198-201
    if (j == g.H) x[c + g.pitch] = (b == 2) ? -raw : raw;
    if (l == 1)   x[c - g.plane] = (b == 3) ? -raw : raw;    // :208-214
    if (l == g.D) x[c + g.plane] = (b == 3) ? -raw : raw;
}

inline int fl_lex_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                        float a, float c)
{
    if (h->nranks > 1)
        return fl_fail(h, FLUID_E_STATE, "FLUID_SOLVER_LEX is a single-GPU parity mode (no z-slabs)");
    const Grid& g = h->g;
    const float c_recip = 1.0f / c;
    LexBatch lb;
    for (int i = 0; i < nb; ++i) { lb.x[i] = h->field[fid[i]]; lb.rhs[i] = h->field[pid[i]]; lb.b[i] = b[i]; }
    const int K = h->acc, smin = 3, smax = g.W + g.H + g.D;
    const int nthreads = g.H * g.D;
    for (int t = smin; t <= smax + 2 * (K - 1); ++t) {
        int kmin = (t - smax + 1) / 2; if (kmin < 0) kmin = 0;      // s = t-2k <= smax
        int kmax = (t - smin) / 2;     if (kmax > K - 1) kmax = K - 1;
        if (kmax < kmin) continue;
        dim3 grid((nthreads + 255) / 256, kmax - kmin + 1, nb);
        k_lex_wavefront<<<grid, 256, 0, h->stream>>>(g, lb, h->cflags, a, c_recip, t, kmin);
        FL_CHECK_LAUNCH(h);
    }
    QuadLaunch q = fl_quad_launch(g, g.Dl);
    for (int i = 0; i < nb; ++i) {
        k_set_bounds<<<q.grid, q.block, 0, h->stream>>>(g, lb.x[i], h->cflags, b[i]);
        FL_CHECK_LAUNCH(h);
    }
    return 0;
}
