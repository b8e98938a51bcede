// sweep_pair_kernel.cuh -- TWO red-black sweeps of linearSolver (simulation.cpp:258-272: relax every
// interior cell, setBounds, twice) in ONE streaming pass over HBM: multi-iteration temporal blocking.
// x in, rhs in, x out once per PAIR of sweeps = 6 B per cell and sweep instead of 12.
//
// Decomposition.  A CTA spans the full x extent (one warp = one pair of rows, a thread = EIGHT
// consecutive cells of two adjacent rows, 136 <= W <= 256, W % 8 == 0) by TR = 2*NP thread rows of
// which TY = 2*NP - 6 are owned (three redundantly computed halo rows per side, one more read-only row
// in the TMA box), and streams along z.  Every thread keeps its z-column in REGISTERS: six planes
//   N  = plane p+1 (old values)          C0 = plane p    C1 = p-1    C2 = p-2    C3 = p-3    C4 = p-4
// and one pipeline step runs four half-sweeps ("stages"), each one plane behind the previous:
//   stage 0: colour 0 of sweep 1 on plane p      stage 1: colour 1 of sweep 1 on plane p-1 (+ setBounds)
//   stage 2: colour 0 of sweep 2 on plane p-2    stage 3: colour 1 of sweep 2 on plane p-3, store
// Colour c of a cell = (x + y + z) & 1, colour 0 first -- the ordering-matched CPU loop of the oracle.
// A cell's update does not read its own value (simulation.cpp:263-269), and in one step all four stages
// act on the same in-register positions (row A: positions i = E mod 2, row B: the others; E alternates
// with p), so the ring just shifts by one plane per step.
//
// Neighbours: x from the own octet + one warp shuffle; y from the pair partner in registers and, for the
// outer row, from shared memory: stage 0 reads the TMA slot of plane p (natural layout), stages 1-3 read
// what the row neighbour published one step earlier -- only the freshly updated colour, stored COMPRESSED
// (4 floats per lane, one conflict-free 128-bit access); z from the register ring.  One __syncthreads per
// step.  The right-hand side of a plane stays in its TMA slot for the four steps that use it.
//
// setBounds between the two sweeps (simulation.cpp:183-246) never touches memory: a ghost cell is read
// only by the one interior cell next to it, so its new value (the neighbour's PRE-zero sweep-1 value,
// negated for the wall-normal velocity component) is kept beside that cell -- the x ghosts in a 4-deep
// register ring (lane 0 / the last lane), the y ghost rows and z ghost planes in the ring entries of
// those rows / planes, which are otherwise idle.  Solid and no-slip cells are zeroed in registers after
// stage 1 (colour 1; the colour-0 values of sweep 1 are never read again) and by the store epilogue.
//
// This file is the DEVICE part only, and is also compiled for the host by tests/host_emul/pair_emul.cpp
// (SP_HOST_EMUL: threads, barriers, shuffles and TMA emulated), so the statements below are checked
// against the CPU oracle in the CPU test suite as well.
#pragma once

#include "sweep_common.cuh"
#ifndef SP_HOST_EMUL
#include "sweep_stream.cuh"              // the PTX wrappers (mbarrier, TMA, dependent launch); the host build emulates them
#endif

#define SP_NX 3                          // x ring: planes p, p+1 (landing), p+2 (in flight)
#define SP_NR 6                          // rhs ring (= mbarriers): planes p-3 .. p+2
#define SP_SROW 128                      // floats per compressed colour row (4 per lane)

template <int NP>
struct SpGeom {
    static constexpr int TR = 2 * NP;                                // thread rows
    static constexpr int TY = 2 * NP - 6;                            // owned rows
    static constexpr int BR = 2 * NP + 2;                            // x box rows
    static constexpr int BOXF = (SW_BOXW * BR + 31) / 32 * 32;       // floats per x box (128-byte multiple)
    static constexpr int XSLOTF = 2 * BOXF;                          // two 136-wide boxes side by side
    static constexpr int RBOXF = SW_RBOXW * TR;
    static constexpr int RSLOTF = 2 * RBOXF;
    static constexpr int SBUFF = (TR + 2) * SP_SROW;                 // one published colour plane (+2 edge rows)
    static constexpr int THREADS = 32 * NP;
    static constexpr int TX_BYTES = 2 * (SW_BOXW * BR + SW_RBOXW * TR) * 4;
    static constexpr size_t SMEM = (size_t)(SP_NX * XSLOTF + SP_NR * RSLOTF + 6 * SBUFF) * sizeof(float);
    static_assert(SMEM <= SW_SMEM_MAX, "pair-sweep tile too large for shared memory");
};

struct alignas(64) PairArgs {
    CUtensorMap map[4];                  // source field of each batched solve (box SW_BOXW x BR x 1)
    CUtensorMap rmap[4];                 // its right-hand side (box SW_RBOXW x TR x 1)
    float* dst[4];
    int b[4];
    const unsigned short* zbits[2];      // [0]: solid, [1]: solid | no-slip; one bit per cell, see k_build_zbits
    const SwSeg* segs;                   // CTA k works through segs[seg_first[k] .. seg_first[k+1])
    const int* seg_first;
    Grid g;
    float a, c_recip;
    int nty;                             // row tiles per field
    int zb_pairs;                        // row pairs per plane of the zbits volume
};

// one plane of a thread's two rows: cells x0 .. x0+7 of row A (a0, a1) and of row B (b0, b1)
struct SpPlane { float4 a0, a1, b0, b1; };

// the 4 cells at positions PE, PE+2, PE+4, PE+6 of an octet
template <int PE>
__device__ __forceinline__ float4 sp_pick(const float4& v0, const float4& v1)
{
    return PE == 0 ? make_float4(v0.x, v0.z, v1.x, v1.z) : make_float4(v0.y, v0.w, v1.y, v1.w);
}
template <int PE>
__device__ __forceinline__ void sp_put(float4& v0, float4& v1, const float4& n)
{
    if (PE == 0) { v0.x = n.x; v0.z = n.y; v1.x = n.z; v1.z = n.w; }
    else         { v0.y = n.x; v0.w = n.y; v1.y = n.z; v1.w = n.w; }
}
__device__ __forceinline__ float4 sp_neg(const float4& v) { return make_float4(-v.x, -v.y, -v.z, -v.w); }

// relax the four cells of one colour of an octet (c0, c1): the x neighbours are the other colour of the
// same octet plus L (left of position 0) or R (right of position 7); yp, ym, zp, zm, r are compressed
template <int PE>
__device__ __forceinline__ float4 sp_relax(const float4& c0, const float4& c1, float L, float R, const float4& yp,
                                           const float4& ym, const float4& zp, const float4& zm, const float4& r,
                                           float a, float cr)
{
    float4 n;
    if (PE == 0) {
        n.x = fl_relax(r.x, a, cr, c0.y, L,    yp.x, ym.x, zp.x, zm.x);
        n.y = fl_relax(r.y, a, cr, c0.w, c0.y, yp.y, ym.y, zp.y, zm.y);
        n.z = fl_relax(r.z, a, cr, c1.y, c0.w, yp.z, ym.z, zp.z, zm.z);
        n.w = fl_relax(r.w, a, cr, c1.w, c1.y, yp.w, ym.w, zp.w, zm.w);
    } else {
        n.x = fl_relax(r.x, a, cr, c0.z, c0.x, yp.x, ym.x, zp.x, zm.x);
        n.y = fl_relax(r.y, a, cr, c1.x, c0.z, yp.y, ym.y, zp.y, zm.y);
        n.z = fl_relax(r.z, a, cr, c1.z, c1.x, yp.z, ym.z, zp.z, zm.z);
        n.w = fl_relax(r.w, a, cr, R,    c1.z, yp.w, ym.w, zp.w, zm.w);
    }
    return n;
}

// zero the cells of colour PE whose bit is set (bits: bit i = position i of the octet)
template <int PE>
__device__ __forceinline__ float4 sp_zero(const float4& n, unsigned bits)
{
    float4 v;
    v.x = (bits & (1u << (PE + 0))) ? 0.0f : n.x;
    v.y = (bits & (1u << (PE + 2))) ? 0.0f : n.y;
    v.z = (bits & (1u << (PE + 4))) ? 0.0f : n.z;
    v.w = (bits & (1u << (PE + 6))) ? 0.0f : n.w;
    return v;
}

// per-thread constants and the small rings that travel with the planes
struct SpThread {
    int lane, last;                      // last = lane that holds x = W
    int offA, roffA, sA;                 // row A: offsets into an x slot, a rhs slot, a published colour plane
    bool intA, intB;                     // row is an interior row of the grid (takes part in the stages)
    bool ghostA, ghostB;                 // row is the y = 0 / y = H+1 ghost row (its pair partner is y = 1 / y = H)
    bool ownA, ownB;                     // row is stored by this CTA
    int smaxA, smaxB;                    // last stage that is still needed on the row (halo rows: fewer)
    bool lane_in;
    unsigned eA, eB;                     // x / y edge bits of the store epilogue
    float xgA[4], xgB[4];                // x ghost beside the row, planes p .. p-3 (lane 0: x = 0, last lane: x = W+1)
    unsigned fz[4];                      // zero bits of planes p .. p-3: row A in bits 0-7, row B in bits 8-15
};

// ---- one half-sweep on one plane ------------------------------------------------------------------
// S = stage, E = parity of the step, UP = planes streamed upwards in z.  c = the plane the stage acts on,
// ahead / behind = its stream neighbours (z+1 / z-1 when UP).  ynbr: S == 0: the x slot of the plane
// (natural layout), else the colour plane published by the previous stage one step ago.  spub: where this
// stage publishes its colour (S < 3).  rslot: the plane's right-hand side.  k = ring position of the plane
// (xg / fz index) = S.
template <int S, int E, bool UP>
__device__ __forceinline__ void sp_stage(SpThread& t, const PairArgs& A, SpPlane& c, SpPlane& ahead, SpPlane& behind,
                                         const float* __restrict__ ynbr, float* __restrict__ spub,
                                         const float* __restrict__ rslot, int b, int zg, bool in_window,
                                         float* __restrict__ out, unsigned oA, bool store_plane)
{
    constexpr int PA = E, PB = 1 - E;                      // active positions of row A / row B
    const float a = A.a, cr = A.c_recip;
    const bool doA = in_window && t.intA && S <= t.smaxA, doB = in_window && t.intB && S <= t.smaxB;
    if (!doA && !doB) return;                              // warp-uniform
    SpPlane& zpP = UP ? ahead : behind;
    SpPlane& zmP = UP ? behind : ahead;
    // x neighbours across lanes: row A needs L when PA == 0, else R; row B the other one
    float LA = 0.0f, RA = 0.0f, LB = 0.0f, RB = 0.0f;
    if (PA == 0) {
        LA = __shfl_up_sync(0xffffffffu, c.a1.w, 1);   if (t.lane == 0) LA = t.xgA[S];
        RB = __shfl_down_sync(0xffffffffu, c.b0.x, 1); if (t.lane == t.last) RB = t.xgB[S];
    } else {
        RA = __shfl_down_sync(0xffffffffu, c.a0.x, 1); if (t.lane == t.last) RA = t.xgA[S];
        LB = __shfl_up_sync(0xffffffffu, c.b1.w, 1);   if (t.lane == 0) LB = t.xgB[S];
    }
    float4 nA = make_float4(0.0f, 0.0f, 0.0f, 0.0f), nB = nA;
    if (doA) {
        float4 ym;
        if (S == 0) ym = sp_pick<PA>(sw_lds4(ynbr + t.offA - SW_BOXW), sw_lds4(ynbr + t.offA - SW_BOXW + 4));
        else        ym = sw_lds4(ynbr + t.sA - SP_SROW);
        const float4 r = sp_pick<PA>(sw_lds4(rslot + t.roffA), sw_lds4(rslot + t.roffA + 4));
        nA = sp_relax<PA>(c.a0, c.a1, LA, RA, sp_pick<PA>(c.b0, c.b1), ym, sp_pick<PA>(zpP.a0, zpP.a1),
                          sp_pick<PA>(zmP.a0, zmP.a1), r, a, cr);
    }
    if (doB) {
        float4 yp;
        if (S == 0) yp = sp_pick<PB>(sw_lds4(ynbr + t.offA + 2 * SW_BOXW), sw_lds4(ynbr + t.offA + 2 * SW_BOXW + 4));
        else        yp = sw_lds4(ynbr + t.sA + 2 * SP_SROW);
        const float4 r = sp_pick<PB>(sw_lds4(rslot + t.roffA + SW_RBOXW), sw_lds4(rslot + t.roffA + SW_RBOXW + 4));
        nB = sp_relax<PB>(c.b0, c.b1, LB, RB, yp, sp_pick<PB>(c.a0, c.a1), sp_pick<PB>(zpP.b0, zpP.b1),
                          sp_pick<PB>(zmP.b0, zmP.b1), r, a, cr);
    }
    // both rows were relaxed from the values before this stage; now they take their new values
    if (S < 2) {
        // ---- sweep 1: what setBounds will make of the ghosts next to these cells (simulation.cpp:187-215),
        // from the PRE-zero values; each ghost is read only by the cell it mirrors ----------------------
        if (doA) {
            if (PA == 0) { if (t.lane == 0) t.xgA[S] = (b == 1) ? -nA.x : nA.x; }        // x = 0      :189
            else         { if (t.lane == t.last) t.xgA[S] = nA.w; }                      // x = W + 1  :191
        }
        if (doB) {
            if (PB == 0) { if (t.lane == 0) t.xgB[S] = (b == 1) ? -nB.x : nB.x; }
            else         { if (t.lane == t.last) t.xgB[S] = nB.w; }
        }
        if (t.ghostA && doB) sp_put<PB>(c.a0, c.a1, b == 2 ? sp_neg(nB) : nB);           // y = 0      :198
        if (t.ghostB && doA) sp_put<PA>(c.b0, c.b1, b == 2 ? sp_neg(nA) : nA);           // y = H + 1  :200
        if (zg == 1) {                                                                   // z = 0      :208
            if (doA) sp_put<PA>(zmP.a0, zmP.a1, b == 3 ? sp_neg(nA) : nA);
            if (doB) sp_put<PB>(zmP.b0, zmP.b1, b == 3 ? sp_neg(nB) : nB);
        }
        if (zg == A.g.D) {                                                               // z = D + 1  :212
            if (doA) sp_put<PA>(zpP.a0, zpP.a1, b == 3 ? sp_neg(nA) : nA);
            if (doB) sp_put<PB>(zpP.b0, zpP.b1, b == 3 ? sp_neg(nB) : nB);
        }
    }
    if (S == 1) {
        // sweep 1 is complete on this plane: solids (and, for velocity components, cells that touch one)
        // become 0 (:219-245).  Only colour 1 needs it: the colour-0 values of sweep 1 are not read again.
        nA = sp_zero<PA>(nA, t.fz[1]);
        nB = sp_zero<PB>(nB, t.fz[1] >> 8);
    }
    if (doA) sp_put<PA>(c.a0, c.a1, nA);
    if (doB) sp_put<PB>(c.b0, c.b1, nB);
    if (S < 3) {
        if (doA) sw_sts4(spub + t.sA, nA);
        if (doB) sw_sts4(spub + t.sA + SP_SROW, nB);
    } else if (store_plane) {
        // ---- store with the setBounds epilogue of sweep 2 (sw_store: ghosts take the pre-zero value) ----
        const Grid& g = A.g;
        const unsigned zbits = (zg == 1 ? SW_ZLO : 0u) | (zg == g.D ? SW_ZHI : 0u);
        const unsigned zmask4 = 0x01010101u;               // the bits below are already class-selected
        if (t.ownA && t.lane_in) {
            const unsigned f = t.fz[3];
            const unsigned lo = (f & 1u) | ((f & 2u) << 7) | ((f & 4u) << 14) | ((f & 8u) << 21);
            const unsigned hi = ((f >> 4) & 1u) | ((f & 0x20u) << 3) | ((f & 0x40u) << 10) | ((f & 0x80u) << 17);
            sw_store(g, out, oA, c.a0, lo, b, zmask4, (t.eA & ~SW_XHI) | zbits);
            sw_store(g, out, oA + 4, c.a1, hi, b, zmask4, (t.eA & ~SW_XLO) | zbits);
        }
        if (t.ownB && t.lane_in) {
            const unsigned f = t.fz[3] >> 8;
            const unsigned lo = (f & 1u) | ((f & 2u) << 7) | ((f & 4u) << 14) | ((f & 8u) << 21);
            const unsigned hi = ((f >> 4) & 1u) | ((f & 0x20u) << 3) | ((f & 0x40u) << 10) | ((f & 0x80u) << 17);
            sw_store(g, out, oA + g.pitch, c.b0, lo, b, zmask4, (t.eB & ~SW_XHI) | zbits);
            sw_store(g, out, oA + g.pitch + 4, c.b1, hi, b, zmask4, (t.eB & ~SW_XLO) | zbits);
        }
    }
}

// ---- one segment: planes za .. zb of one column -----------------------------------------------------
struct SpRing {                         // TMA ring position, runs on from segment to segment
    int xs_iss, rs_iss;                  // x slot / rhs slot (= barrier) of the next plane issued
    int xs_con, rs_con; unsigned ph_con; // ... of the next plane consumed, and its mbarrier parity
};

template <int NP, bool UP>
__device__ __forceinline__ void sp_segment(const PairArgs& A, const SwSeg sg, SpThread& t, float* const xring,
                                           float* const rring, float* const sbuf, unsigned long long* const bars,
                                           SpRing& R, const int tid, const int w)
{
    using GE = SpGeom<NP>;
    constexpr int XSLOTF = GE::XSLOTF, RSLOTF = GE::RSLOTF, BOXF = GE::BOXF, RBOXF = GE::RBOXF, SBUFF = GE::SBUFF;
    constexpr int DIR = UP ? 1 : -1;
    const Grid& g = A.g;
    const int batch = sg.col / A.nty;
    const int y0 = 1 + (sg.col - batch * A.nty) * GE::TY;        // first owned row (odd)
    const int za = sg.za, zb = sg.zb;
    const int b = A.b[batch];
    float* __restrict__ out = A.dst[batch];
    const unsigned short* __restrict__ zbv = A.zbits[b == 0 ? 0 : 1];

    // windows: stage s runs on planes za-(3-s) .. zb+(3-s), clipped to the interior 1 .. D
    const int wlo0 = max(za - 3, 1), whi0 = min(zb + 3, g.D);
    const int wlo1 = max(za - 2, 1), whi1 = min(zb + 2, g.D);
    const int wlo2 = max(za - 1, 1), whi2 = min(zb + 1, g.D);
    const int ps = UP ? wlo0 : whi0;                             // first stage-0 plane
    const int pe = UP ? zb + 3 : za - 3;                         // plane index of the last step (stage 3 on zb / za)
    const int nsteps = (pe - ps) * DIR + 1;
    const int first = ps - DIR;                                  // first plane loaded (0 .. D+1 on one GPU)
    const int lastp = UP ? min(zb + 4, g.D + 1) : max(za - 4, 0);
    const int nplanes = (lastp - first) * DIR + 1;

    auto issue = [&](int p) {                                    // plane p -> next ring slots
        if (tid == 0) {
            unsigned long long* bar = &bars[R.rs_iss];
            sw_mbar_expect_tx(bar, GE::TX_BYTES);
            float* xs = xring + R.xs_iss * XSLOTF;
            float* rs = rring + R.rs_iss * RSLOTF;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                // the tensor maps start at the allocation's plane -1, hence p + 1
                sw_tma_load_3d(xs + i * BOXF, &A.map[batch], 128 * i, y0 - 4, p + 1, bar);
                sw_tma_load_3d(rs + i * RBOXF, &A.rmap[batch], FL_XOFF + 1 + 128 * i, y0 - 3, p + 1, bar);
            }
        }
        if (++R.xs_iss == SP_NX) R.xs_iss = 0;
        if (++R.rs_iss == SP_NR) R.rs_iss = 0;
    };
    int issued = min(nplanes, SP_NX);
    for (int j = 0; j < issued; ++j) issue(first + DIR * j);

    // this thread's rows
    const int yA = y0 - 3 + 2 * w, yB = yA + 1;
    const int x0 = 1 + 8 * t.lane;
    t.intA = yA >= 1 && yA <= g.H; t.intB = yB >= 1 && yB <= g.H;
    t.ghostA = yA == 0; t.ghostB = yB == g.H + 1;
    t.ownA = t.intA && w >= 2 && w <= NP - 2;
    t.ownB = t.intB && w >= 1 && w <= NP - 3;
    t.smaxA = min(2 * w, GE::TR - 1 - 2 * w); t.smaxB = min(2 * w + 1, GE::TR - 2 - 2 * w);
    t.eA = (yA == 1 ? SW_YLO : 0u) | (yA == g.H ? SW_YHI : 0u) | (x0 == 1 ? SW_XLO : 0u) | (x0 + 7 == g.W ? SW_XHI : 0u);
    t.eB = (yB == 1 ? SW_YLO : 0u) | (yB == g.H ? SW_YHI : 0u) | (x0 == 1 ? SW_XLO : 0u) | (x0 + 7 == g.W ? SW_XHI : 0u);
    // element offset of row A's octet in plane 0 (32-bit: the padded volume has < 2^32 cells); rows that are
    // not stored never use it
    const unsigned oA0 = (unsigned)fl_idx(g, x0, yA >= 0 ? yA : 0, 0);
    const unsigned plane = (unsigned)g.plane;
    const unsigned zb_row = (unsigned)((yA >> 1) * 32 + t.lane);  // yA is even; pairs (2k, 2k+1)
    const bool zb_ok = yA >= 0 && yA <= g.H;

    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    SpPlane N, C0, C1, C2, C3, C4;
    N.a0 = N.a1 = N.b0 = N.b1 = zero4; C0 = N; C1 = N; C2 = N; C3 = N; C4 = N;
#pragma unroll
    for (int i = 0; i < 4; ++i) { t.xgA[i] = 0.0f; t.xgB[i] = 0.0f; t.fz[i] = 0u; }

    // slot bookkeeping: rs0 = rhs slot of plane p (stage s reads slot rs0 - s), xs0 = x slot of plane p
    auto consume = [&](const float*& xs, int& rs) {              // wait for the next plane in stream order
        sw_mbar_wait(&bars[R.rs_con], R.ph_con);
        xs = xring + R.xs_con * XSLOTF; rs = R.rs_con;
        if (++R.xs_con == SP_NX) R.xs_con = 0;
        if (++R.rs_con == SP_NR) { R.rs_con = 0; R.ph_con ^= 1u; }
    };
    auto load_plane = [&](SpPlane& P, const float* xs) {
        P.a0 = sw_lds4(xs + t.offA);           P.a1 = sw_lds4(xs + t.offA + 4);
        P.b0 = sw_lds4(xs + t.offA + SW_BOXW); P.b1 = sw_lds4(xs + t.offA + SW_BOXW + 4);
    };
    const float* xs_p; int rs0;
    { const float* xb; int rb; consume(xb, rb); load_plane(C1, xb); }        // the plane behind ps
    consume(xs_p, rs0); load_plane(C0, xs_p);                                // plane ps
    __syncthreads();        // the slot of the plane behind ps is re-armed by the first step

    int p = ps, jc = 1;                                          // jc = stream index of plane p
    // one pipeline step with static parity EE = (z_global(p) + 1) & 1; the published colour planes are
    // double-buffered by that parity (written [EE], read [1 - EE])
#define SP_STEP(EE)                                                                                             \
    {                                                                                                           \
        if (issued < nplanes && issued - jc < SP_NX) { issue(first + DIR * issued); ++issued; }                 \
        const bool has_next = jc + 1 < nplanes;                                                                 \
        const float* xs_n = xs_p; int rs_n = rs0 + 1 == SP_NR ? 0 : rs0 + 1;    /* the ring position moves on even past the last plane */ \
        if (has_next) { consume(xs_n, rs_n); load_plane(N, xs_n); }                                             \
        /* x ghosts beside plane p as the last setBounds left them, and the plane's zero bits */               \
        t.xgA[0] = xs_p[t.offA + (t.lane == 0 ? -1 : 8)];                                                       \
        t.xgB[0] = xs_p[t.offA + SW_BOXW + (t.lane == 0 ? -1 : 8)];                                             \
        t.fz[0] = (zb_ok && p >= 1 && p <= g.D) ? (unsigned)zbv[(unsigned)p * (unsigned)(A.zb_pairs * 32) + zb_row] : 0u; \
        const int zg = g.zlo + p - 1;                                                                           \
        float* const S0w = sbuf + (0 + (EE)) * SBUFF;     const float* const S0r = sbuf + (0 + 1 - (EE)) * SBUFF; \
        float* const S1w = sbuf + (2 + (EE)) * SBUFF;     const float* const S1r = sbuf + (2 + 1 - (EE)) * SBUFF; \
        float* const S2w = sbuf + (4 + (EE)) * SBUFF;     const float* const S2r = sbuf + (4 + 1 - (EE)) * SBUFF; \
        const int q1 = p - DIR, q2 = p - 2 * DIR, q3 = p - 3 * DIR;                                             \
        sp_stage<0, EE, UP>(t, A, C0, N, C1, xs_p, S0w, rring + rs0 * RSLOTF, b, zg,                            \
                            p >= wlo0 && p <= whi0, out, 0u, false);                                            \
        sp_stage<1, EE, UP>(t, A, C1, C0, C2, S0r, S1w, rring + ((rs0 + SP_NR - 1) % SP_NR) * RSLOTF, b, zg - DIR, \
                            q1 >= wlo1 && q1 <= whi1, out, 0u, false);                                          \
        sp_stage<2, EE, UP>(t, A, C2, C1, C3, S1r, S2w, rring + ((rs0 + SP_NR - 2) % SP_NR) * RSLOTF, b, zg - 2 * DIR, \
                            q2 >= wlo2 && q2 <= whi2, out, 0u, false);                                          \
        sp_stage<3, EE, UP>(t, A, C3, C2, C4, S2r, nullptr, rring + ((rs0 + SP_NR - 3) % SP_NR) * RSLOTF, b, zg - 3 * DIR, \
                            q3 >= za && q3 <= zb, out, oA0 + (unsigned)q3 * plane, true);                       \
        C4 = C3; C3 = C2; C2 = C1; C1 = C0; C0 = N;                                                             \
        t.xgA[3] = t.xgA[2]; t.xgA[2] = t.xgA[1]; t.xgA[1] = t.xgA[0];                                          \
        t.xgB[3] = t.xgB[2]; t.xgB[2] = t.xgB[1]; t.xgB[1] = t.xgB[0];                                          \
        t.fz[3] = t.fz[2]; t.fz[2] = t.fz[1]; t.fz[1] = t.fz[0];                                                \
        xs_p = xs_n; rs0 = rs_n; p += DIR; ++jc;                                                                \
        __syncthreads();                                                                                        \
    }
    int step = 0;
    if (((g.zlo + ps) & 1) == 0) {                               // EE of the first step = (zg + 1) & 1 with zg = zlo + ps - 1
        for (;;) { if (step++ == nsteps) break; SP_STEP(0) if (step++ == nsteps) break; SP_STEP(1) }
    } else {
        for (;;) { if (step++ == nsteps) break; SP_STEP(1) if (step++ == nsteps) break; SP_STEP(0) }
    }
#undef SP_STEP
}

// One launch = two red-black sweeps of up to four batched fields.  A CTA works through a short list of
// segments (column = batched field x row tile, planes za .. zb).  sp_cta is the whole CTA program; the
// kernel below hands it the shared memory (the host emulation calls it from one host thread per CUDA thread).
template <int NP, bool UP>
__device__ __forceinline__ void sp_cta(const PairArgs& A, float* const sw_mem, unsigned long long* const bars,
                                       const int tid, const int bid)
{
    using GE = SpGeom<NP>;
    float* const xring = sw_mem;
    float* const rring = sw_mem + SP_NX * GE::XSLOTF;
    float* const sbuf = rring + SP_NR * GE::RSLOTF;

    const int lane = tid & 31, w = tid >> 5;
    if (tid == 0) {
        for (int i = 0; i < SP_NR; ++i) sw_mbar_init(&bars[i], 1);
        sw_fence_mbar_init();
    }
    sw_grid_dep_sync();                  // programmatic dependent launch, see sweep_stream.cuh
    __syncthreads();

    SpThread t;
    t.lane = lane; t.last = A.g.W / 8 - 1;
    t.lane_in = lane <= t.last;
    t.offA = (lane >> 4) * GE::BOXF + (1 + 2 * w) * SW_BOXW + 4 + 8 * (lane & 15);
    t.roffA = (lane >> 4) * GE::RBOXF + (2 * w) * SW_RBOXW + 8 * (lane & 15);
    t.sA = (2 * w + 1) * SP_SROW + 4 * lane;

    SpRing R; R.xs_iss = R.rs_iss = R.xs_con = R.rs_con = 0; R.ph_con = 0u;
    for (int si = A.seg_first[bid]; si < A.seg_first[bid + 1]; ++si)
        sp_segment<NP, UP>(A, A.segs[si], t, xring, rring, sbuf, bars, R, tid, w);
}

#ifndef SP_HOST_EMUL
template <int NP, bool UP>
__global__ void __launch_bounds__(32 * NP, 1) k_sweep_pair(const __grid_constant__ PairArgs A)
{
    extern __shared__ __align__(128) float sw_mem[];
    __shared__ unsigned long long bars[SP_NR];
    sp_cta<NP, UP>(A, sw_mem, bars, (int)threadIdx.x, (int)blockIdx.x);
}
#endif
