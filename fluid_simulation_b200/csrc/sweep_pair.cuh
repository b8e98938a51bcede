// sweep_pair.cuh -- host side of the temporally blocked sweep (two red-black sweeps per HBM pass,
// sweep_pair_kernel.cuh): zero-bit volume, launch sequence of a solve.  One GPU, red-black,
// 136 <= W <= 256, W % 8 == 0, H even; everything else stays with k_sweep_stream (one sweep per pass).
#pragma once
#include "sweep_pair_kernel.cuh"

#define SP_SEGCOST 8                     // extra pipeline steps per segment (6 fill / drain + prologue)

// FLUID_PAIR=1 switches the temporally blocked kernel on.  It is OFF by default: measured on B200 at 256^3 it
// halves the HBM traffic per sweep (140 instead of 242 MB, ncu) but is bound by instruction issue at 8 warps
// per SM and ends up level with k_sweep_stream (47.7 vs 46.5 us per single-field sweep, 44.6 vs 42.9 us in the
// step mix; profiles/r02_pair_*).  Read per solve so that tests can A/B inside one process.
inline bool fl_pair_eligible(fluid_handle* h)
{
    const char* on = getenv("FLUID_PAIR");
    const Grid& g = h->g;
    if (!on || atoi(on) == 0 || h->nranks != 1 || (h->flags & FLUID_SOLVER_MASK) != FLUID_SOLVER_RB) return false;
    return g.W > 128 && g.W <= 256 && (g.W & 7) == 0 && (g.H & 1) == 0 && g.cells < (1ll << 32);
}

template <int NP, int NU>
inline int sw_pair_run(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid, float a, float c,
                       unsigned x_from_rhs, int npairs)
{
    using GE = SpGeom<NP>;
    const Grid& g = h->g;
    StreamState* s = sw_state(h);
    typedef void (*pair_kernel)(const PairArgs);
    const pair_kernel kerns[2] = { k_sweep_pair<NP, true, NU>, k_sweep_pair<NP, false, NU> };
    for (int i = 0; i < 2; ++i)
        FL_CUDA(h, cudaFuncSetAttribute(kerns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GE::SMEM));
    if (nb > 4) return fl_fail(h, FLUID_E_ARG, "streaming solve: %d batched fields", nb);
    SwPlan plan;
    if (int e = sw_get_plan(h, nb, GE::TY, 1, 0, &plan, SP_SEGCOST)) return e;

    if (int e = sw_get_zbits(h)) return e;                                // zero bits, rebuilt whenever the flag volume was

    PairArgs A;
    A.g = g; A.a = a; A.c_recip = 1.0f / c;                              // simulation.cpp:257
    A.nty = (g.H + GE::TY - 1) / GE::TY;
    A.segs = plan.segs; A.seg_first = plan.seg_first;
    A.zbits[0] = s->zbits[0]; A.zbits[1] = s->zbits[1]; A.zb_pairs = s->zb_pairs;   // zb_lanes == 32 (W <= 256)

    float* cur[4]; float* alt[4];
    if (nb == 4 && !h->scratch[3]) {                                      // fourth ping-pong partner, first use
        const size_t fb = (size_t)g.cells * sizeof(float);
        FL_CUDA(h, cudaMalloc(&h->scratch[3], fb));
        FL_CUDA(h, cudaMemsetAsync(h->scratch[3], 0, fb, h->stream));
        h->scratch[3] += FL_ZPAD * g.plane;
    }
    for (int i = 0; i < nb; ++i) {
        cur[i] = h->field[fid[i]]; alt[i] = h->scratch[i];
        A.b[i] = b[i];
        if (int e = sw_get_map(h, h->field[pid[i]], SW_RBOXW, GE::TR, &A.rmap[i])) return e;
    }
    dim3 ge((2 * (g.W + 2) + 2 * g.H + 127) / 128, g.Dl + 2);
    for (int i = 0; i < nb; ++i) {                                        // edge/corner ghosts follow (see sw_run)
        if ((x_from_rhs >> i) & 1u) {
            k_copy_edges<<<ge, 128, 0, h->stream>>>(g, h->field[pid[i]], cur[i]);
            FL_CHECK_LAUNCH(h);
            k_copy_edges<<<ge, 128, 0, h->stream>>>(g, h->field[pid[i]], alt[i]);
        } else
            k_copy_edges<<<ge, 128, 0, h->stream>>>(g, cur[i], alt[i]);
        FL_CHECK_LAUNCH(h);
    }
    for (int k = 0; k < npairs; ++k) {                                    // simulation.cpp:258, two iterations at a time
        const unsigned from_rhs = k == 0 ? x_from_rhs : 0u;               // see sw_run
        for (int i = 0; i < nb; ++i) {
            const bool fr = (from_rhs >> i) & 1u;
            if (int e = sw_get_map(h, fr ? h->field[pid[i]] : cur[i], SW_BOXW, GE::BR, &A.map[i])) return e;
            A.dst[i] = fr ? cur[i] : alt[i];
        }
        const pair_kernel kern = kerns[(s->zigzag && (k & 1)) ? 1 : 0];   // odd passes stream the other way (L2 reuse)
        if (s->pdl >= 1) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(plan.nctas); cfg.blockDim = dim3(GE::THREADS); cfg.dynamicSmemBytes = GE::SMEM; cfg.stream = h->stream;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            FL_CUDA(h, cudaLaunchKernelEx(&cfg, kern, A));
        } else
            kern<<<plan.nctas, GE::THREADS, GE::SMEM, h->stream>>>(A);
        FL_CHECK_LAUNCH(h);
        for (int i = 0; i < nb; ++i) if (!((from_rhs >> i) & 1u)) { float* t = cur[i]; cur[i] = alt[i]; alt[i] = t; }
    }
    for (int i = 0; i < nb; ++i)
        if (cur[i] != h->field[fid[i]]) { h->scratch[i] = h->field[fid[i]]; h->field[fid[i]] = cur[i]; }
    return 0;
}

// the first 2 * floor(acc / 2) sweeps of a solve, two per pass; *done = sweeps performed
int fl_pair_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid, float a, float c,
                         unsigned x_from_rhs, int* done)
{
    *done = 0;
    const int npairs = h->acc / 2;
    if (npairs <= 0 || !fl_pair_eligible(h)) return 0;
    // FLUID_PAIR_UNROLL=6: six unrolled steps (no register moves, three times the code; measured slower:
    // instruction fetch) instead of the two-step loop body with a ring rotation
    const char* u = getenv("FLUID_PAIR_UNROLL");
    if (!u || atoi(u) == 2) { if (int e = sw_pair_run<8, 2>(h, nb, b, fid, pid, a, c, x_from_rhs, npairs)) return e; }
    else                   { if (int e = sw_pair_run<8, 6>(h, nb, b, fid, pid, a, c, x_from_rhs, npairs)) return e; }
    *done = 2 * npairs;
    return 0;
}
