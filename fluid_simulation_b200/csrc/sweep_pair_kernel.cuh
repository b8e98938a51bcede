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
// Colour c of a cell = (x + y + z) & 1, colour 0 first -- the order of the CPU checker's red-black loop.
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

#define SP_NX 4                          // x ring: planes p, p+1, p+2 (landing), p+3 (in flight): TMA runs two steps ahead
#define SP_NR (SP_NX + 2)                // rhs ring (= mbarriers): a plane's right-hand side is read at steps p and p+2
#define SP_SROW 128                      // floats per compressed colour row (4 per lane)
// The register ring: six planes in six fixed slots.  In unrolled step U the plane p+1-j (j = 0: the plane being
// loaded, 1: plane p, ..., 5: plane p-4) lives in slot SP_SLOT(j, U), so advancing by one plane renames
// registers instead of moving them; NU steps make one loop body (6: no moves at all; 2: the ring is
// rotated by two slots at the end of the body -- a third of the code).
#define SP_SLOT(j, U) (((j) - (U) + 12) % 6)

template <int NP>
struct SpGeom {
    static constexpr int TR = 2 * NP;                                // thread rows
    static constexpr int TY = 2 * NP - 6;                            // owned rows
    static constexpr int BR = 2 * NP + 2;                            // x box rows
    static constexpr int BOXF = (SW_BOXW * BR + 31) / 32 * 32;       // floats per x box (128-byte multiple)
    static constexpr int XSLOTF = 2 * BOXF;                          // two 136-wide boxes side by side
    static constexpr int RBOXF = SW_RBOXW * TR;
    static constexpr int RSLOTF = 2 * RBOXF;
    static constexpr int SBUFF = TR * SP_SROW;                       // one published colour plane (the outer neighbour rows
                                                                     // of the first / last thread row fall into the adjacent
                                                                     // buffer or the pad row: never-used values)
    static constexpr int THREADS = 32 * NP;
    static constexpr int TX_BYTES = 2 * (SW_BOXW * BR + SW_RBOXW * TR) * 4;
    static constexpr size_t SMEM = (size_t)(SP_NX * XSLOTF + SP_NR * RSLOTF + 6 * SBUFF + SP_SROW) * sizeof(float);
    static_assert(SMEM + 64 <= SW_SMEM_MAX, "pair-sweep tile too large for shared memory");
};

struct alignas(64) PairArgs {
    CUtensorMap map[4];                  // source field of each batched solve (box SW_BOXW x BR x 1)
    CUtensorMap rmap[4];                 // its right-hand side (box SW_RBOXW x TR x 1)
    float* dst[4];
    int b[4];
    const unsigned short* zbits[2];      // [0]: solid, [1]: solid | no-slip; one bit per cell, local planes 1-FL_HALO .. Dl+FL_HALO (k_build_zbits)
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
    bool ownA, ownB;                     // this thread's octet of the row is stored by this CTA
    bool lane_in;
    unsigned eA, eB;                     // x / y edge bits of the store epilogue
    // rings indexed like the plane ring (SP_SLOT): only the entries of planes p .. p-3 are live
    float xgA[6], xgB[6];                // x ghost beside the row (lane 0: x = 0, last lane: x = W+1)
    unsigned fz[6];                      // zero bits: row A in bits 0-7, row B in bits 8-15
    bool lo_lane, hi_lane;               // lane 0 / the lane that holds x = W
    float4 h1A, h1B, h3A, h3B;           // right-hand side of the cells stage 1 / stage 3 relax NEXT step (loaded with the other colour)
    unsigned sink;                       // keeps the unused lanes of 128-bit shared loads alive (see sp_lds_half)
};

// 128-bit shared loads of an octet of which only the colour PE is consumed.  ptxas would narrow them to four
// scalar LDS -- each an 8-way bank conflict across the warp (lane stride 32 B) -- so the unused lanes are
// folded into a sink word (one LOP3 per load).
template <int PE>
__device__ __forceinline__ float4 sp_lds_half(const float* p, unsigned& sink)
{
    const float4 v0 = sw_lds4(p), v1 = sw_lds4(p + 4);
    if (PE == 0) sink |= (__float_as_uint(v0.y) | __float_as_uint(v0.w)) | (__float_as_uint(v1.y) | __float_as_uint(v1.w));
    else         sink |= (__float_as_uint(v0.x) | __float_as_uint(v0.z)) | (__float_as_uint(v1.x) | __float_as_uint(v1.z));
    return sp_pick<PE>(v0, v1);
}

// ---- the rare part of the store epilogue, out of line: face ghosts of one finished octet row ---------
// (simulation.cpp:187-215: ghosts take the PRE-zero value; x = 0 negates for b == 1, y faces for b == 2,
// z faces for b == 3, x = W+1 is a plain copy).  Called by the few threads that sit on a face.
__device__ __noinline__ void sp_store_ghosts(const Grid g, float* __restrict__ out, unsigned c, float4 r0, float4 r1,
                                             int b, unsigned edge)
{
    if (edge & SW_XLO) out[c - 1] = (b == 1) ? -r0.x : r0.x;
    if (edge & SW_XHI) out[c + 8] = r1.w;
    if (edge & (SW_YLO | SW_YHI)) {
        const float4 g0 = (b == 2) ? sp_neg(r0) : r0, g1 = (b == 2) ? sp_neg(r1) : r1;
        if (edge & SW_YLO) { st4(out + c - g.pitch, g0); st4(out + c - g.pitch + 4, g1); }
        if (edge & SW_YHI) { st4(out + c + g.pitch, g0); st4(out + c + g.pitch + 4, g1); }
    }
    if (edge & (SW_ZLO | SW_ZHI)) {
        const float4 g0 = (b == 3) ? sp_neg(r0) : r0, g1 = (b == 3) ? sp_neg(r1) : r1;
        if (edge & SW_ZLO) { st4(out + c - g.plane, g0); st4(out + c - g.plane + 4, g1); }
        if (edge & SW_ZHI) { st4(out + c + g.plane, g0); st4(out + c + g.plane + 4, g1); }
    }
}

// one finished octet row: zero solids / no-slip cells (bits: bit i = position i), store, ghosts if on a face
__device__ __forceinline__ void sp_store_row(const Grid& g, float* __restrict__ out, unsigned c, const float4& r0,
                                             const float4& r1, unsigned bits, int b, unsigned edge)
{
    st4(out + c,     make_float4((bits & 1u) ? 0.0f : r0.x, (bits & 2u) ? 0.0f : r0.y, (bits & 4u) ? 0.0f : r0.z,
                                 (bits & 8u) ? 0.0f : r0.w));
    st4(out + c + 4, make_float4((bits & 16u) ? 0.0f : r1.x, (bits & 32u) ? 0.0f : r1.y, (bits & 64u) ? 0.0f : r1.z,
                                 (bits & 128u) ? 0.0f : r1.w));
    if (edge) sp_store_ghosts(g, out, c, r0, r1, b, edge);
}

// ---- one half-sweep on one plane ------------------------------------------------------------------
// S = stage, U = unroll position (parity E = U & 1), UP = planes streamed upwards in z.  c = the plane the
// stage acts on, ahead / behind = its stream neighbours (z+1 / z-1 when UP).  ynbr: S == 0: the x slot of
// the plane (natural layout), else the colour plane published by the previous stage one step ago.  spub:
// where this stage publishes its colour (S < 3).  rA / rB: right-hand side of the active cells.
// Both rows are relaxed unconditionally in one basic block (eight independent cells in flight): a row that
// lies outside the grid, or is a halo row the stage is no longer needed on, computes values nobody uses;
// only the y ghost rows have to keep theirs (intA / intB guard the write-back).
template <int S, int U, bool UP>
__device__ __forceinline__ void sp_stage(SpThread& t, const PairArgs& A, SpPlane& c, SpPlane& ahead, SpPlane& behind,
                                         const float* __restrict__ ynbr, float* __restrict__ spub,
                                         const float4& rA, const float4& rB, int b, int zg,
                                         float* __restrict__ out, unsigned oA, bool store_plane)
{
    constexpr int E = U & 1;                               // segments start on an even step (see sp_segment)
    constexpr int PA = E, PB = 1 - E;                      // active positions of row A / row B
    constexpr int K = SP_SLOT(S + 1, U), K1 = SP_SLOT(2, U), K3 = SP_SLOT(4, U);   // ring slots of this plane, of planes p-1 / p-3
    const float a = A.a, cr = A.c_recip;
    SpPlane& zpP = UP ? ahead : behind;
    SpPlane& zmP = UP ? behind : ahead;
    // x neighbours across lanes: row A needs L when PA == 0, else R; row B the other one
    float LA = 0.0f, RA = 0.0f, LB = 0.0f, RB = 0.0f;
    if (PA == 0) {
        LA = __shfl_up_sync(0xffffffffu, c.a1.w, 1);   LA = t.lo_lane ? t.xgA[K] : LA;
        RB = __shfl_down_sync(0xffffffffu, c.b0.x, 1); RB = t.hi_lane ? t.xgB[K] : RB;
    } else {
        RA = __shfl_down_sync(0xffffffffu, c.a0.x, 1); RA = t.hi_lane ? t.xgA[K] : RA;
        LB = __shfl_up_sync(0xffffffffu, c.b1.w, 1);   LB = t.lo_lane ? t.xgB[K] : LB;
    }
    float4 ym, yp;
    if (S == 0) { ym = sp_lds_half<PA>(ynbr + t.offA - SW_BOXW, t.sink); yp = sp_lds_half<PB>(ynbr + t.offA + 2 * SW_BOXW, t.sink); }
    else        { ym = sw_lds4(ynbr + t.sA - SP_SROW);                   yp = sw_lds4(ynbr + t.sA + 2 * SP_SROW); }
    float4 nA = sp_relax<PA>(c.a0, c.a1, LA, RA, sp_pick<PA>(c.b0, c.b1), ym, sp_pick<PA>(zpP.a0, zpP.a1),
                             sp_pick<PA>(zmP.a0, zmP.a1), rA, a, cr);
    float4 nB = sp_relax<PB>(c.b0, c.b1, LB, RB, yp, sp_pick<PB>(c.a0, c.a1), sp_pick<PB>(zpP.b0, zpP.b1),
                             sp_pick<PB>(zmP.b0, zmP.b1), rB, a, cr);
    if (S < 2) {
        // ---- sweep 1: what setBounds will make of the ghosts next to these cells (simulation.cpp:187-215),
        // from the PRE-zero values; each ghost is read only by the cell it mirrors ----------------------
        if (PA == 0) {
            t.xgA[K] = t.lo_lane ? ((b == 1) ? -nA.x : nA.x) : t.xgA[K];                 // x = 0      :189
            t.xgB[K] = t.hi_lane ? nB.w : t.xgB[K];                                      // x = W + 1  :191
        } else {
            t.xgA[K] = t.hi_lane ? nA.w : t.xgA[K];
            t.xgB[K] = t.lo_lane ? ((b == 1) ? -nB.x : nB.x) : t.xgB[K];
        }
        if (t.ghostA) sp_put<PB>(c.a0, c.a1, b == 2 ? sp_neg(nB) : nB);                  // y = 0      :198
        if (t.ghostB) sp_put<PA>(c.b0, c.b1, b == 2 ? sp_neg(nA) : nA);                  // y = H + 1  :200
        if (zg == 1) {                                                                   // z = 0      :208
            sp_put<PA>(zmP.a0, zmP.a1, b == 3 ? sp_neg(nA) : nA);
            sp_put<PB>(zmP.b0, zmP.b1, b == 3 ? sp_neg(nB) : nB);
        }
        if (zg == A.g.D) {                                                               // z = D + 1  :212
            sp_put<PA>(zpP.a0, zpP.a1, b == 3 ? sp_neg(nA) : nA);
            sp_put<PB>(zpP.b0, zpP.b1, b == 3 ? sp_neg(nB) : nB);
        }
    }
    if (S == 1) {
        // sweep 1 is complete on this plane: solids (and, for velocity components, cells that touch one)
        // become 0 (:219-245).  Only colour 1 needs it: the colour-0 values of sweep 1 are not read again.
        nA = sp_zero<PA>(nA, t.fz[K1]);
        nB = sp_zero<PB>(nB, t.fz[K1] >> 8);
    }
    if (t.intA) sp_put<PA>(c.a0, c.a1, nA);
    if (t.intB) sp_put<PB>(c.b0, c.b1, nB);
    if (S < 3) {
        sw_sts4(spub + t.sA, nA);
        sw_sts4(spub + t.sA + SP_SROW, nB);
    } else if (store_plane) {
        // ---- store with the setBounds epilogue of sweep 2 ----
        const Grid& g = A.g;
        const unsigned zbits = (zg == 1 ? SW_ZLO : 0u) | (zg == g.D ? SW_ZHI : 0u);
        if (t.ownA) sp_store_row(g, out, oA, c.a0, c.a1, t.fz[K3] & 0xffu, b, t.eA | zbits);
        if (t.ownB) sp_store_row(g, out, oA + (unsigned)g.pitch, c.b0, c.b1, (t.fz[K3] >> 8) & 0xffu, b, t.eB | zbits);
    }
}

// ---- one segment: planes za .. zb of one column -----------------------------------------------------
struct SpRing {                         // TMA ring position, runs on from segment to segment
    int xs_iss, rs_iss;                  // x slot / rhs slot (= barrier) of the next plane issued
    int xs_con, rs_con; unsigned ph_con; // ... of the next plane consumed, and its mbarrier parity
};

template <int NP, bool UP, int NU>
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

    // windows: stage s runs on planes za-(3-s) .. zb+(3-s), clipped to the interior of the TUNNEL (global
    // z = 1 .. D are the local planes zmin .. zmax; a z-slab recomputes up to three planes of each neighbour)
    const int zmin = 2 - g.zlo, zmax = g.D - g.zlo + 1;
    const int wlo0 = max(za - 3, zmin), whi0 = min(zb + 3, zmax);
    const int wlo1 = max(za - 2, zmin), whi1 = min(zb + 2, zmax);
    const int wlo2 = max(za - 1, zmin), whi2 = min(zb + 1, zmax);
    // first stage-0 plane; the unrolled body fixes the parity of a step by its position, so a segment whose
    // first plane has the odd parity starts one (idle) step earlier: every window excludes that plane
    const int ps0 = UP ? wlo0 : whi0;
    const int ps = ((g.zlo + ps0) & 1) ? ps0 - DIR : ps0;
    const int pe = UP ? zb + 3 : za - 3;                         // plane index of the last step (stage 3 on zb / za)
    const int nsteps = (pe - ps) * DIR + 1;
    const int first = ps - DIR;                                  // first plane loaded (-FL_ZPAD .. Dl+1+FL_ZPAD: inside the allocation)
    const int lastp = UP ? min(zb + 4, zmax + 1) : max(za - 4, zmin - 1);
    const int nplanes = (lastp - first) * DIR + 1;

    auto issue = [&](int p) {                                    // plane p -> next ring slots
        if (tid == 0) {
            unsigned long long* bar = &bars[R.rs_iss];
            sw_mbar_expect_tx(bar, GE::TX_BYTES);
            float* xs = xring + R.xs_iss * XSLOTF;
            float* rs = rring + R.rs_iss * RSLOTF;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                // the tensor maps start at the allocation's first plane, -FL_ZPAD
                sw_tma_load_3d(xs + i * BOXF, &A.map[batch], 128 * i, y0 - 4, p + FL_ZPAD, bar);
                sw_tma_load_3d(rs + i * RBOXF, &A.rmap[batch], FL_XOFF + 1 + 128 * i, y0 - 3, p + FL_ZPAD, bar);
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
    t.ownA = t.intA && w >= 2 && w <= NP - 2 && t.lane_in;
    t.ownB = t.intB && w >= 1 && w <= NP - 3 && t.lane_in;
    t.eA = (yA == 1 ? SW_YLO : 0u) | (yA == g.H ? SW_YHI : 0u) | (x0 == 1 ? SW_XLO : 0u) | (x0 + 7 == g.W ? SW_XHI : 0u);
    t.eB = (yB == 1 ? SW_YLO : 0u) | (yB == g.H ? SW_YHI : 0u) | (x0 == 1 ? SW_XLO : 0u) | (x0 + 7 == g.W ? SW_XHI : 0u);
    // element offset of row A's octet in plane 0 (32-bit: the padded volume has < 2^32 cells); rows that are
    // not stored never use it
    const unsigned oA0 = (unsigned)fl_idx(g, x0, yA >= 0 ? yA : 0, 0);
    const unsigned plane = (unsigned)g.plane;
    const unsigned zb_row = (unsigned)((yA >> 1) * 32 + t.lane);  // yA is even; pairs (2k, 2k+1)
    const bool zb_ok = yA >= 0 && yA <= g.H;
    const int zlo_bits = max(zmin, 1 - FL_HALO), zhi_bits = min(zmax, g.Dl + FL_HALO);   // planes the zero-bit volume covers

    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    SpPlane P[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) { P[i].a0 = P[i].a1 = P[i].b0 = P[i].b1 = zero4; t.xgA[i] = 0.0f; t.xgB[i] = 0.0f; t.fz[i] = 0u; }

    // slot bookkeeping: xs_p = x slot of plane p; rs0 / rs1 / rs2 = rhs slots of planes p, p-1, p-2
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
    const float* xs_p; int rs0, rs1, rs2;
    { const float* xb; int rb; consume(xb, rb); load_plane(P[SP_SLOT(2, 0)], xb); }   // the plane behind ps
    consume(xs_p, rs0); load_plane(P[SP_SLOT(1, 0)], xs_p);                           // plane ps
    rs1 = rs2 = rs0;
    t.h1A = t.h1B = t.h3A = t.h3B = zero4;
    __syncthreads();        // the slot of the plane behind ps is re-armed by the first step

    const bool edge_lane = t.lo_lane || t.hi_lane;
    const int xg_off = t.offA + (t.lo_lane ? -1 : 8);            // the x ghost beside row A in an x slot
    int p = ps, jc = 1;                                          // jc = stream index of plane p
    // One pipeline step at unroll position U (parity U & 1 = (z_global(p) + 1) & 1); the published colour
    // planes are double-buffered by that parity (written [E], read [1 - E]).  The right-hand side of a plane is
    // read twice, as whole octets: at step p (colour 0 of sweep 1 now, the other colour kept for stage 1 of the
    // next step) and at step p+2 (the same for sweep 2), so every lane of every 128-bit load is used.
#define SP_STEP(U)                                                                                              \
    {                                                                                                           \
        constexpr int EE = (U) & 1;                                                                             \
        SpPlane& N = P[SP_SLOT(0, U)]; SpPlane& C0 = P[SP_SLOT(1, U)]; SpPlane& C1 = P[SP_SLOT(2, U)];          \
        SpPlane& C2 = P[SP_SLOT(3, U)]; SpPlane& C3 = P[SP_SLOT(4, U)]; SpPlane& C4 = P[SP_SLOT(5, U)];         \
        if (issued < nplanes && issued - jc < SP_NX) { issue(first + DIR * issued); ++issued; }                 \
        const bool has_next = jc + 1 < nplanes;                                                                 \
        const float* xs_n = xs_p; int rs_n = rs0 + 1 == SP_NR ? 0 : rs0 + 1;    /* the ring position moves on even past the last plane */ \
        if (has_next) { consume(xs_n, rs_n); load_plane(N, xs_n); }                                             \
        /* x ghosts beside plane p as the last setBounds left them, and the plane's zero bits */               \
        t.xgA[SP_SLOT(1, U)] = edge_lane ? xs_p[xg_off] : 0.0f;                                                 \
        t.xgB[SP_SLOT(1, U)] = edge_lane ? xs_p[xg_off + SW_BOXW] : 0.0f;                                       \
        t.fz[SP_SLOT(1, U)] = (zb_ok && p >= zlo_bits && p <= zhi_bits) ? (unsigned)zbv[(unsigned)(p + FL_HALO - 1) * (unsigned)(A.zb_pairs * 32) + zb_row] : 0u; \
        const int zg = g.zlo + p - 1;                                                                           \
        float* const S0w = sbuf + (0 + EE) * SBUFF;     const float* const S0r = sbuf + (0 + 1 - EE) * SBUFF;   \
        float* const S1w = sbuf + (2 + EE) * SBUFF;     const float* const S1r = sbuf + (2 + 1 - EE) * SBUFF;   \
        float* const S2w = sbuf + (4 + EE) * SBUFF;     const float* const S2r = sbuf + (4 + 1 - EE) * SBUFF;   \
        const int q1 = p - DIR, q2 = p - 2 * DIR, q3 = p - 3 * DIR;                                             \
        {                                                                                                       \
            const float* rp = rring + rs0 * RSLOTF + t.roffA;                                                   \
            const float4 u0 = sw_lds4(rp), u1 = sw_lds4(rp + 4), u2 = sw_lds4(rp + SW_RBOXW), u3 = sw_lds4(rp + SW_RBOXW + 4); \
            if (p >= wlo0 && p <= whi0)                                                                         \
                sp_stage<0, U, UP>(t, A, C0, N, C1, xs_p, S0w, sp_pick<EE>(u0, u1), sp_pick<1 - EE>(u2, u3), b, zg, out, 0u, false); \
            if (q1 >= wlo1 && q1 <= whi1)                                                                       \
                sp_stage<1, U, UP>(t, A, C1, C0, C2, S0r, S1w, t.h1A, t.h1B, b, zg - DIR, out, 0u, false);      \
            t.h1A = sp_pick<1 - EE>(u0, u1); t.h1B = sp_pick<EE>(u2, u3);                                       \
        }                                                                                                       \
        {                                                                                                       \
            const float* rp = rring + rs2 * RSLOTF + t.roffA;                                                   \
            const float4 u0 = sw_lds4(rp), u1 = sw_lds4(rp + 4), u2 = sw_lds4(rp + SW_RBOXW), u3 = sw_lds4(rp + SW_RBOXW + 4); \
            if (q2 >= wlo2 && q2 <= whi2)                                                                       \
                sp_stage<2, U, UP>(t, A, C2, C1, C3, S1r, S2w, sp_pick<EE>(u0, u1), sp_pick<1 - EE>(u2, u3), b, zg - 2 * DIR, out, 0u, false); \
            if (q3 >= za && q3 <= zb)                                                                           \
                sp_stage<3, U, UP>(t, A, C3, C2, C4, S2r, nullptr, t.h3A, t.h3B, b, zg - 3 * DIR, out, oA0 + (unsigned)q3 * plane, true); \
            t.h3A = sp_pick<1 - EE>(u0, u1); t.h3B = sp_pick<EE>(u2, u3);                                       \
        }                                                                                                       \
        xs_p = xs_n; rs2 = rs1; rs1 = rs0; rs0 = rs_n; p += DIR; ++jc;                                          \
        __syncthreads();                                                                                        \
    }
    int step = 0;
    for (;;) {
        if (step++ == nsteps) break;
        SP_STEP(0)
        if (step++ == nsteps) break;
        SP_STEP(1)
        if constexpr (NU == 6) {
            if (step++ == nsteps) break;
            SP_STEP(2)
            if (step++ == nsteps) break;
            SP_STEP(3)
            if (step++ == nsteps) break;
            SP_STEP(4)
            if (step++ == nsteps) break;
            SP_STEP(5)
        } else {   // two steps done: rotate the rings by two slots so that the body starts at position 0 again
            SpPlane Q[6]; float qa[6], qb[6]; unsigned qf[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) { Q[i] = P[i]; qa[i] = t.xgA[i]; qb[i] = t.xgB[i]; qf[i] = t.fz[i]; }
#pragma unroll
            for (int i = 0; i < 6; ++i) { P[i] = Q[(i + 4) % 6]; t.xgA[i] = qa[(i + 4) % 6]; t.xgB[i] = qb[(i + 4) % 6]; t.fz[i] = qf[(i + 4) % 6]; }
        }
    }
#undef SP_STEP
}

// One launch = two red-black sweeps of up to four batched fields.  A CTA works through a short list of
// segments (column = batched field x row tile, planes za .. zb).  sp_cta is the whole CTA program; the
// kernel below hands it the shared memory (the host emulation calls it from one host thread per CUDA thread).
template <int NP, bool UP, int NU>
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
    t.sA = (2 * w) * SP_SROW + 4 * lane;
    t.lo_lane = lane == 0; t.hi_lane = lane == t.last; t.sink = 0u;

    SpRing R; R.xs_iss = R.rs_iss = R.xs_con = R.rs_con = 0; R.ph_con = 0u;
    for (int si = A.seg_first[bid]; si < A.seg_first[bid + 1]; ++si)
        sp_segment<NP, UP, NU>(A, A.segs[si], t, xring, rring, sbuf, bars, R, tid, w);
    if (A.nty < 0) A.dst[0][0] = __uint_as_float(t.sink);            // never taken: keeps t.sink observable
}

#ifndef SP_HOST_EMUL
template <int NP, bool UP, int NU>
__global__ void __launch_bounds__(32 * NP, 1) k_sweep_pair(const __grid_constant__ PairArgs A)
{
    extern __shared__ __align__(128) float sw_mem[];
    __shared__ unsigned long long bars[SP_NR];
    sp_cta<NP, UP, NU>(A, sw_mem, bars, (int)threadIdx.x, (int)blockIdx.x);
}
#endif
