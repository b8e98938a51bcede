// sweep_stream.cuh -- the linearSolver sweep (simulation.cpp:259-271: relax every interior cell,
// then setBounds) as ONE streaming pass over HBM per sweep: x in, rhs in, x out = the 12 B/cell
// algorithmic minimum (+1 B/cell of flags).
//
// Decomposition.  A CTA spans the full x extent of the grid (TXW warps of 32 x-quads = 128 cells
// each, W <= 128*TXW) by TY = 2*NP-2 owned rows (+1 halo row above and below), and streams along z
// through a chunk of planes.  Every thread owns one x-quad of two adjacent rows and keeps its own
// z-column in REGISTERS; only the y-neighbour rows and the two x-edge scalars of a warp come from
// shared memory.
//   * plane p+1 of the source field arrives through TMA (cp.async.bulk.tensor 3-D box per warp
//     column, 136 floats x BR rows, zero fill outside the array) into a ring of SW_NSLOT slots,
//     SW_AHEAD planes ahead of the compute, signalled on one mbarrier per slot;
//   * rhs and the flag bytes are read straight from global memory into registers one plane ahead
//     (coalesced 16-byte loads, each byte used once);
//   * red-black: iteration p first relaxes the colour-0 cells of plane p (owned rows + halo rows)
//     from the old values in the TMA slots and registers ("stage 0"), publishes that plane to a
//     double-buffered shared-memory slot for the row neighbours, then relaxes the colour-1 cells of
//     plane p-1 from colour-0 values only ("stage 1") and stores the finished quad through the
//     fused setBounds epilogue.  One __syncthreads per plane.  With the quad origin x0 = 1 (mod 4)
//     and the first row of a pair even, both stages of an iteration touch the same two lanes of a
//     quad: lanes (E, E+2) in the even row and (1-E, 3-E) in the odd row, E = (z_global(p)+1) & 1.
//   * Jacobi: stage 0 is the identity and stage 1 relaxes all four lanes.
// Solid cells are relaxed like any other cell and only zeroed by the epilogue of the sweep, so
// colour-1 cells see the transient colour-0 value of a solid neighbour exactly as the ordering-
// matched CPU loop does.  Arithmetic is fl_relax (no FMA, reference operand order): bit-identical
// to the simple kernels and to the CPU oracle.
#pragma once
#include "handle.cuh"
#include "kernels_stencil.cuh"
#include "sweep_common.cuh"
#include <cuda.h>
#include <map>
#include <cstdlib>

struct alignas(64) StreamArgs {
    CUtensorMap map[4];                  // source field of each batched solve (box SW_BOXW x BR x 1)
    CUtensorMap rmap[4];                 // its right-hand side (box SW_RBOXW x BR-2 x 1)
    float* dst[4];
    int b[4];
    const unsigned short* zbits[2];      // zero bits, one per cell ([0]: solid, [1]: solid | no-slip), see k_build_zbits
    unsigned zb_plane; int zb_lanes;     // words per plane / per row pair of that volume
    const SwSeg* segs;                   // CTA k works through segs[seg_first[k] .. seg_first[k+1])
    const int* seg_first;
    Grid g;
    float a, c_recip;
    int nty;                             // row tiles per field
    int nxt, xt_cells;                   // x tiles per row tile and the cells each owns (W > 128 * TXW: a CTA spans part of x,
                                         // one redundantly computed halo quad beside every interior tile edge); 1, W otherwise
    int flip;                            // 1 = this sweep streams every segment against its planned direction
    // z-slabs, halos handed over inside the sweep (segments with a face in SwSeg.mode): the neighbour's
    // mapping of dst[] per face, its arrival flag for this rank, this rank's per-face CTA counters
    float* peer_dst[2][4];
    unsigned* peer_flag[2];
    unsigned* counter;
    int peer_zdst[2];                    // neighbour-local index of the first of the two halo planes written
    int n_edge;                          // edge segments per face (= columns)
    unsigned seq;                        // value the flags are raised to
    const unsigned* my_flag;             // this rank's arrival flags [face]; an edge segment starts once its face shows seq - 1
    unsigned* err_word;                  // set when an arrival flag did not show up within SW_ARRIVAL_TIMEOUT_NS
    int dbg;                             // FLUID_SW_DBG timing experiments: 1 = no stores, 2 = no relaxation (results invalid)
    unsigned long long* trace;           // FLUID_SW_TRACE: per CTA { start, first segment done, end } in globaltimer ns; else nullptr
};

// ---- PTX wrappers (mbarrier + TMA) -----------------------------------------------------------
__device__ __forceinline__ unsigned sw_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void sw_mbar_init(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sw_smem(bar)), "r"(count));
}
__device__ __forceinline__ void sw_mbar_expect_tx(unsigned long long* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sw_smem(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void sw_mbar_wait(unsigned long long* bar, unsigned parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n"
        "W_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra W_%=;\n\t}"
        ::"r"(sw_smem(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void sw_tma_load_3d(void* dst, const CUtensorMap* map, int c0, int c1, int c2,
                                               unsigned long long* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(sw_smem(dst)), "l"(reinterpret_cast<unsigned long long>(map)), "r"(c0), "r"(c1), "r"(c2),
          "r"(sw_smem(bar)) : "memory");
}
__device__ __forceinline__ void sw_fence_mbar_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// Programmatic dependent launch (no-ops unless the launch carries the attribute): the next sweep of the
// solve may be scheduled while this one drains; no field data is touched before the kernel this one
// depends on has completed and flushed.
__device__ __forceinline__ void sw_grid_dep_sync()
{
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void sw_fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// cell left of a quad (L) / right of it (R): the neighbour lane's outer cell by shuffle; the two
// edge lanes of the warp read the halo column of the shared-memory box instead.
__device__ __forceinline__ float sw_left(const float4& c, const float* row_quad, int lane)
{
    float L = __shfl_up_sync(0xffffffffu, c.w, 1);
    if (lane == 0) L = row_quad[-1];
    return L;
}
__device__ __forceinline__ float sw_right(const float4& c, const float* row_quad, int lane)
{
    float R = __shfl_down_sync(0xffffffffu, c.x, 1);
    if (lane == 31) R = row_quad[4];
    return R;
}

// relax the two lanes (E, E+2) of quad c in place.  L / R are the cells left of lane 0 / right of
// lane 3; only the one that the colour needs is read.
template <int E>
__device__ __forceinline__ void sw_relax2(float4& c, float L, float R, const float4& yp, const float4& ym,
                                          const float4& zp, const float4& zm, const float4& r,
                                          float a, float cr)
{
    if (E == 0) {
        const float n0 = fl_relax(r.x, a, cr, c.y, L,   yp.x, ym.x, zp.x, zm.x);
        const float n2 = fl_relax(r.z, a, cr, c.w, c.y, yp.z, ym.z, zp.z, zm.z);
        c.x = n0; c.z = n2;
    } else {
        const float n1 = fl_relax(r.y, a, cr, c.z, c.x, yp.y, ym.y, zp.y, zm.y);
        const float n3 = fl_relax(r.w, a, cr, R,   c.z, yp.w, ym.w, zp.w, zm.w);
        c.y = n1; c.w = n3;
    }
}

template <int TXW, int NP, int CPS = 1>
struct SwGeom {
    static constexpr int BR = 2 * NP + 2;                            // x box rows: TY owned + 2 halo + 2 read-only
    static constexpr int TY = 2 * NP - 2;                            // owned rows per CTA
    static constexpr int BOXF = (SW_BOXW * BR + 31) / 32 * 32;       // floats per x box, 128-byte multiple
    static constexpr int RBOXF = SW_RBOXW * (BR - 2);                // floats per rhs box (owned + halo rows)
    static constexpr int XSLOTF = TXW * BOXF;                        // the x boxes of one plane
    static constexpr int SLOTF = TXW * (BOXF + RBOXF);               // one plane of the CTA: x boxes, then rhs boxes
    static constexpr int THREADS = 32 * TXW * NP;
    static constexpr int TX_BYTES = TXW * (SW_BOXW * BR + SW_RBOXW * (BR - 2)) * 4;
    // ring depth: as many planes in flight as shared memory allows (the TMA ring IS the latency hiding:
    // one CTA per SM, so the bytes in flight are the bytes of the slots being filled)
    static constexpr int SMEM_BUDGET = CPS == 1 ? SW_SMEM_MAX : (SW_SM_SMEM / CPS - 2048);      // per CTA
    static constexpr int NSLOT_FIT = (SMEM_BUDGET - 2 * XSLOTF * 4 - 1024) / (SLOTF * 4);
    static constexpr int NSLOT = NSLOT_FIT > 8 ? 8 : NSLOT_FIT;
    static constexpr int AHEAD = NSLOT - 2;                          // TMA prefetch distance in planes
    static constexpr size_t SMEM = (size_t)(NSLOT * SLOTF + 2 * XSLOTF) * sizeof(float);
    static_assert(NSLOT >= 3, "tile too large for the shared-memory ring");
};

// Per-thread state of the streaming sweep: the register-resident z-column of the two rows and
// everything that is constant along z.  The column is a ring of four planes: in unrolled step U
// Q[U] is plane p-2 (two behind), Q[U+1] plane p-1, Q[U+2] plane p and Q[U+3] receives plane p+1,
// so advancing by one plane renames registers instead of moving them.  R (rhs, read from the TMA
// slot of plane p) and F (flag bytes, fetched from global memory three planes ahead of their use)
// follow the same indexing.
struct SwThread {
    float4 QA[4], QB[4], RA[4], RB[4];
    unsigned FA[4];                        // zero-bit words (row A: bits 0-7, row B: bits 8-15 of the thread's octet)
    const unsigned short* zbv; unsigned zoff; int zsh;   // this segment's bit volume, the thread's word in a plane, its nibble
    unsigned oA, oB;                       // element offsets of the two quads in plane p+1
    int offA, offB;                        // shared-memory offsets of the two x quads inside a slot
    int roffA, roffB;                      // ... and of the two rhs quads
    int lane;
    unsigned eA, eB;                       // x / y edge bits
    unsigned sink;                         // keeps the dead lanes of 128-bit shared loads alive (see sw_lds4)
    bool intA, intB, ownA, ownB, pub_gl, pub_gr, pub_nl, pub_nr;
};

// 128-bit shared load of a row neighbour of which only lanes (E, E+2) are consumed.  ptxas would
// narrow it to two scalar LDS -- each a 4-way bank conflict across the warp (lane stride 16 B), 8
// wavefronts instead of 4 -- so the two dead lanes are folded into a sink word (one LOP3).
template <int E>
__device__ __forceinline__ float4 sw_lds4_half(const float* p, unsigned& sink)
{
    const float4 v = sw_lds4(p);
    if (E == 0) sink |= __float_as_uint(v.y) | __float_as_uint(v.w);
    else        sink |= __float_as_uint(v.x) | __float_as_uint(v.z);
    return v;
}

// one plane of the pipeline: stage 0 on plane p, stage 1 on the plane one step behind (Jacobi: the
// whole update of plane p).  U = position in the 4-step unrolled loop, E = (z_global(p) + 1) & 1 and
// DIR = +1 / -1 (planes streamed upwards / downwards) are static.  "next" / "behind" below are in
// stream order; zp / zm of the relaxation are picked from them according to DIR so the operand order
// of the reference (z+1 before z-1) is kept in both directions.
template <int BOXF, int XSLOTF, int MODE, int DIR, int U, int E>
__device__ __forceinline__ void sw_step(SwThread& t, const StreamArgs& A, float* __restrict__ out, int b,
                                        int p, int za, int zb, bool has_next,
                                        const float* Up, const float* Un, unsigned long long* bar_n,
                                        unsigned phase_n, float* vbuf)
{
    const Grid& g = A.g;
    const float a = A.a, cr = A.c_recip;
    const unsigned dplane = DIR > 0 ? (unsigned)g.plane : 0u - (unsigned)g.plane;   // one step, mod 2^32
    constexpr int I2 = U & 3, IM = (U + 1) & 3, IC = (U + 2) & 3, IN = (U + 3) & 3;
    const int offA = t.offA, offB = t.offB;
    // flag bytes two steps ahead (t.oA addresses the next plane): consumed by the store of that
    // plane, which is three steps away for red-black and two for Jacobi
    if (p + 2 * DIR >= za && p + 2 * DIR <= zb) {
        t.FA[I2] = t.zbv[(unsigned)(p + 2 * DIR + FL_HALO - 1) * A.zb_plane + t.zoff];
    }
    // Every shared-memory operand of the step is fetched up front, in one burst, so that the LDS
    // latency is exposed once per plane instead of once per stage: rhs and the row neighbours of plane
    // p (TMA slot), the row neighbours of the plane behind (published last step), then the own column
    // of the next plane as soon as its TMA has landed.
    t.RA[IC] = sw_lds4(Up + t.roffA); t.RB[IC] = sw_lds4(Up + t.roffB);
    float4& cA = t.QA[IC]; float4& cB = t.QB[IC];
    const float4& mA = t.QA[IM]; const float4& mB = t.QB[IM];
    const float* Vq = vbuf + ((U + 1) & 1) * XSLOTF;                 // published (and synced) last step
    const int q = p - DIR;
    // stage 0 runs on every plane that is interior to the TUNNEL: on a z-slab that includes the first
    // halo plane, whose colour-0 update is recomputed locally from the second halo plane
    const bool do0 = MODE == 1 && (unsigned)(g.zlo + p - 2) < (unsigned)g.D && !(A.dbg & 2), do1 = MODE == 1 && q >= za && q <= zb && !(A.dbg & 2);
    float4 ymA0 = cA, ypB0 = cB, ymA1 = cA, ypB1 = cB;
    float L0 = 0.0f, R0 = 0.0f, L1 = 0.0f, R1 = 0.0f;
    if (do0) {
        ymA0 = sw_lds4_half<E>(Up + offA - SW_BOXW, t.sink);
        ypB0 = sw_lds4_half<1 - E>(Up + offB + SW_BOXW, t.sink);
        if (E == 0) { L0 = sw_left(cA, Up + offA, t.lane); R0 = sw_right(cB, Up + offB, t.lane); }
        else        { R0 = sw_right(cA, Up + offA, t.lane); L0 = sw_left(cB, Up + offB, t.lane); }
    }
    if (do1) {
        ymA1 = sw_lds4_half<E>(Vq + offA - SW_BOXW, t.sink);
        ypB1 = sw_lds4_half<1 - E>(Vq + offB + SW_BOXW, t.sink);
        if (E == 0) { L1 = sw_left(mA, Vq + offA, t.lane); R1 = sw_right(mB, Vq + offB, t.lane); }
        else        { R1 = sw_right(mA, Vq + offA, t.lane); L1 = sw_left(mB, Vq + offB, t.lane); }
    }
    if (has_next) {
        sw_mbar_wait(bar_n, phase_n);
        t.QA[IN] = sw_lds4(Un + offA); t.QB[IN] = sw_lds4(Un + offB);
    }
    // stage 0 / Jacobi at plane p: z+1 is the next plane when streaming upwards
    const float4& zpA0 = DIR > 0 ? t.QA[IN] : mA; const float4& zmA0 = DIR > 0 ? mA : t.QA[IN];
    const float4& zpB0 = DIR > 0 ? t.QB[IN] : mB; const float4& zmB0 = DIR > 0 ? mB : t.QB[IN];

    if (MODE == 0) {
        // ---- Jacobi: relax all four lanes of plane p from the old planes p-1, p, p+1 --------------
        if (p >= za && p <= zb) {
            const float4 ymA = sw_lds4(Up + offA - SW_BOXW), ypB = sw_lds4(Up + offB + SW_BOXW);
            const float LA = sw_left(cA, Up + offA, t.lane), RA = sw_right(cA, Up + offA, t.lane);
            const float LB = sw_left(cB, Up + offB, t.lane), RB = sw_right(cB, Up + offB, t.lane);
            const unsigned zbits = (g.zlo + p - 1 == 1 ? SW_ZLO : 0u) | (g.zlo + p - 1 == g.D ? SW_ZHI : 0u);
            float4 e = cA, f = cA;
            sw_relax2<0>(e, LA, RA, cB, ymA, zpA0, zmA0, t.RA[IC], a, cr);
            sw_relax2<1>(f, LA, RA, cB, ymA, zpA0, zmA0, t.RA[IC], a, cr);
            if (t.ownA) sw_store_bits(g, out, t.oA - dplane, make_float4(e.x, f.y, e.z, f.w), t.FA[IC] >> t.zsh, b, t.eA | zbits);
            e = cB; f = cB;
            sw_relax2<0>(e, LB, RB, ypB, cA, zpB0, zmB0, t.RB[IC], a, cr);
            sw_relax2<1>(f, LB, RB, ypB, cA, zpB0, zmB0, t.RB[IC], a, cr);
            if (t.ownB) sw_store_bits(g, out, t.oB - dplane, make_float4(e.x, f.y, e.z, f.w), t.FA[IC] >> (8 + t.zsh), b, t.eB | zbits);
        }
    } else {
        // ---- stage 0: colour-0 cells of plane p, in registers --------------------------------------
        if (do0) {
            float4 tA = cA, tB = cB;
            if (E == 0) {
                sw_relax2<0>(tA, L0, 0.0f, cB, ymA0, zpA0, zmA0, t.RA[IC], a, cr);
                sw_relax2<1>(tB, 0.0f, R0, ypB0, cA, zpB0, zmB0, t.RB[IC], a, cr);
            } else {
                sw_relax2<1>(tA, 0.0f, R0, cB, ymA0, zpA0, zmA0, t.RA[IC], a, cr);
                sw_relax2<0>(tB, L0, 0.0f, ypB0, cA, zpB0, zmB0, t.RB[IC], a, cr);
            }
            if (t.intA) cA = tA;
            if (t.intB) cB = tB;
        }
        // publish plane p (colour 0 new, colour 1 old) for the row / warp neighbours
        float* Vp = vbuf + (U & 1) * XSLOTF;
        sw_sts4(Vp + offA, cA); sw_sts4(Vp + offB, cB);
        if (t.pub_nl) { Vp[offA + 128 - BOXF] = cA.x; Vp[offB + 128 - BOXF] = cB.x; }
        if (t.pub_nr) { Vp[offA + BOXF - 125] = cA.w; Vp[offB + BOXF - 125] = cB.w; }
        if (t.pub_gl) { Vp[offA - 1] = Up[offA - 1]; Vp[offB - 1] = Up[offB - 1]; }           // x = 0 ghosts
        if (t.pub_gr) { Vp[offA + 4] = Up[offA + 4]; Vp[offB + 4] = Up[offB + 4]; }           // x = W+1 ghosts
        // ---- stage 1: colour-1 cells of the plane one step behind, from colour-0 values; setBounds --
        if (do1) {
            // around plane q: plane p is one step ahead of it, ring entry I2 one step behind it
            const float4& zpA1 = DIR > 0 ? cA : t.QA[I2]; const float4& zmA1 = DIR > 0 ? t.QA[I2] : cA;
            const float4& zpB1 = DIR > 0 ? cB : t.QB[I2]; const float4& zmB1 = DIR > 0 ? t.QB[I2] : cB;
            float4 tA = mA, tB = mB;
            if (E == 0) {
                sw_relax2<0>(tA, L1, 0.0f, mB, ymA1, zpA1, zmA1, t.RA[IM], a, cr);
                sw_relax2<1>(tB, 0.0f, R1, ypB1, mA, zpB1, zmB1, t.RB[IM], a, cr);
            } else {
                sw_relax2<1>(tA, 0.0f, R1, mB, ymA1, zpA1, zmA1, t.RA[IM], a, cr);
                sw_relax2<0>(tB, L1, 0.0f, ypB1, mA, zpB1, zmB1, t.RB[IM], a, cr);
            }
            const int gzq = g.zlo + q - 1;
            const unsigned zbits = (gzq == 1 ? SW_ZLO : 0u) | (gzq == g.D ? SW_ZHI : 0u);
            if (t.ownA && !(A.dbg & 1)) sw_store_bits(g, out, t.oA - 2u * dplane, tA, t.FA[IM] >> t.zsh, b, t.eA | zbits);
            if (t.ownB && !(A.dbg & 1)) sw_store_bits(g, out, t.oB - 2u * dplane, tB, t.FA[IM] >> (8 + t.zsh), b, t.eB | zbits);
        }
    }
}

// An edge segment reads the z-neighbour's halo planes of the previous sweep and overwrites the
// neighbour's halo planes of the sweep before that: both are safe once the neighbour's flag shows the
// previous sweep (its edge CTAs of that face are done).  The flag is raised by a kernel on ANOTHER GPU
// that does not depend on this launch; normally it was raised long ago.  The wait is bounded: a missing
// neighbour becomes an error word the host reports (fluid_sync), not a hung GPU.
#define SW_ARRIVAL_TIMEOUT_NS 20000000000ull
__device__ __forceinline__ void sw_await_face(const StreamArgs& A, int side, int tid)
{
    if (tid == 0) {
        const unsigned want = A.seq - 1u;
        unsigned v;
        unsigned long long t0 = 0;
        for (unsigned spins = 0;; ++spins) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(A.my_flag + side) : "memory");
            if ((int)(v - want) >= 0) break;
            if ((spins & 1023u) == 1023u) {
                unsigned long long t;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                if (t0 == 0) t0 = t;
                else if (t - t0 > SW_ARRIVAL_TIMEOUT_NS) { atomicExch(A.err_word, 1u + (unsigned)side); break; }
            }
        }
        asm volatile("fence.proxy.async;" ::: "memory");              // the halo planes are read by TMA
    }
    __syncthreads();
}

// The two boundary planes of a face segment are final: copy this column's part of them (its owned rows;
// the first / last row tile also takes the y ghost rows) into the z-neighbour's halo planes over NVLink.
// Nothing waits for the stores here; sw_signal_face fences them when the segment is done.
template <int TY, int THREADS>
__device__ __forceinline__ void sw_push_face(const StreamArgs& A, const Grid& g, const float* out, int batch, int tile,
                                             int y0, int q_lo, int q_hi, int side, int tid)
{
    // (the __syncthreads that ended the previous step made the CTA's stores visible to the CTA)
    // rows r0 .. r1, float4 columns q_lo .. q_hi-1 of each row: the cells this column's CTA wrote (an x tile
    // takes only its own part of a row; the first / last tile also the x ghosts and the row padding)
    const int r0 = tile == 0 ? 0 : y0;
    const int r1 = tile == A.nty - 1 ? g.rows - 1 : y0 + TY - 1;
    const int nq = q_hi - q_lo, n4 = (r1 - r0 + 1) * nq, p4 = g.pitch / 4;
    const long long zsrc = side == 0 ? 1 : g.Dl - 1;
#pragma unroll
    for (int pl = 0; pl < 2; ++pl) {
        const float4* __restrict__ s4 = reinterpret_cast<const float4*>(out + (zsrc + pl) * g.plane + (long long)r0 * g.pitch);
        float4* __restrict__ d4 = reinterpret_cast<float4*>(A.peer_dst[side][batch] + ((long long)A.peer_zdst[side] + pl) * g.plane +
                                                            (long long)r0 * g.pitch);
        for (int i = tid; i < n4; i += THREADS) { const int r = i / nq, q = q_lo + (i - r * nq); d4[r * p4 + q] = __ldcg(s4 + r * p4 + q); }
    }
}

// The face segment is done: fence the hand-over stores (long since on their way) and let the last CTA of
// the face raise the neighbour's arrival flag.
__device__ __forceinline__ void sw_signal_face(const StreamArgs& A, int side, int tid)
{
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
        const unsigned done = atomicAdd(A.counter + side, 1u);
        if (done == (unsigned)A.n_edge - 1u) {                        // every CTA of this face has fenced its stores
            A.counter[side] = 0;
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(A.peer_flag[side]), "r"(A.seq) : "memory");
        }
    }
}

// One segment (planes za..zb of one column) streamed in direction DIR (+1 upwards in z, -1 downwards);
// the TMA ring position and its mbarrier phase (s_iss, s_con, ph_con) run on from segment to segment.
template <int TXW, int NP, int CPS, int MODE, int DIR>
__device__ __forceinline__ void sw_segment(const StreamArgs& A, const SwSeg sg, SwThread& t, float* const uring,
                                           float* const vbuf, unsigned long long* const bars, int& s_iss, int& s_con,
                                           unsigned& ph_con, const int tid, const int pr, const int rA, const int qx,
                                           const unsigned safe)
{
    using GE = SwGeom<TXW, NP, CPS>;
    constexpr int BOXF = GE::BOXF, RBOXF = GE::RBOXF, SLOTF = GE::SLOTF, XSLOTF = GE::XSLOTF;
    constexpr int NSLOT = GE::NSLOT;
    const Grid& g = A.g;
    const unsigned plane = (unsigned)g.plane;
    const unsigned dplane = DIR > 0 ? plane : 0u - plane;
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    // column = ((batched field * x tiles) + x tile) * row tiles + row tile
    const int bx = sg.col / A.nty;
    const int batch = bx / A.nxt, xtile = bx - batch * A.nxt;
    const int y0 = 1 + (sg.col - bx * A.nty) * GE::TY;           // first owned row (odd)
    const int own_lo = 1 + xtile * A.xt_cells;                   // first cell the tile owns
    const int own_hi = min(own_lo + A.xt_cells - 1, g.W);
    const int xo = own_lo - 1 - (xtile > 0 ? 4 : 0);             // x offset of the CTA's first quad (a halo quad inside the grid)
    const int x0 = 1 + xo + 4 * qx;                              // this thread's quad: cells x0 .. x0+3
    const bool lane_in = x0 <= g.W;
    const bool own_x = x0 >= own_lo && x0 <= own_hi;
    const unsigned exbits = (x0 == 1 ? SW_XLO : 0u) | (x0 + 3 == g.W ? SW_XHI : 0u);
    const int za = sg.za, zb = sg.zb;
    const int lo = max(za - 2, g.zl_lo), hi = min(zb + 2, g.zl_hi);   // planes this segment reads
    const int ps = DIR > 0 ? za - 1 : zb + 1;                    // first stage-0 plane
    const int nsteps = zb - za + 3;
    const bool has_behind = DIR > 0 ? (ps - 1 >= lo) : (ps + 1 <= hi);
    const int first = has_behind ? ps - DIR : ps;                // first plane loaded, then first + DIR, ...
    const int nplanes = (DIR > 0 ? hi - first : first - lo) + 1;
    const int b = A.b[batch];
    float* __restrict__ out = A.dst[batch];
    if ((sg.mode >> 1) && !(A.dbg & 4)) sw_await_face(A, (sg.mode >> 1) - 1, tid);     // dbg 4: timing experiment, results invalid

    // plane p -> ring slot s_iss.  Every thread tracks the ring position, so that the slot and barrier
    // addresses stay warp-uniform; thread 0 alone arms the barrier and starts the copies.
    auto issue = [&](int p) {
        if (tid == 0) {
            unsigned long long* bar = &bars[s_iss];
            sw_mbar_expect_tx(bar, GE::TX_BYTES);
            float* slot = uring + s_iss * SLOTF;
#pragma unroll
            for (int i = 0; i < TXW; ++i) {
                // the tensor maps start at the allocation's first plane, -FL_ZPAD
                sw_tma_load_3d(slot + i * BOXF, &A.map[batch], xo + 128 * i, y0 - 2, p + FL_ZPAD, bar);
                sw_tma_load_3d(slot + XSLOTF + i * RBOXF, &A.rmap[batch], FL_XOFF + 1 + xo + 128 * i, y0 - 1, p + FL_ZPAD, bar);
            }
        }
        if (++s_iss == NSLOT) s_iss = 0;
    };
    int issued = min(nplanes, NSLOT);                            // planes handed to TMA so far (uniform)
    for (int j = 0; j < issued; ++j) issue(first + DIR * j);

    // this thread's quad: cells x0..x0+3 of rows yA (even) and yB = yA + 1 (odd)
    const int yA = y0 - 2 + rA, yB = yA + 1;
    t.intA = lane_in && yA >= 1 && yA <= g.H;                    // row takes part in stage 0
    t.intB = lane_in && yB >= 1 && yB <= g.H;
    t.ownA = t.intA && pr >= 1 && own_x;                         // quad is stored by this CTA
    t.ownB = t.intB && pr <= NP - 2 && own_x;
    t.eA = exbits | (yA == 1 ? SW_YLO : 0u) | (yA == g.H ? SW_YHI : 0u);
    t.eB = exbits | (yB == 1 ? SW_YLO : 0u) | (yB == g.H ? SW_YHI : 0u);
    // element offsets (32-bit: the padded volume has < 2^32 cells).  Rows / lanes outside the
    // grid read a harmless in-range cell instead of being predicated off.
    t.oA = (t.intA ? (unsigned)fl_idx(g, x0, yA, 0) : safe) + (unsigned)ps * plane;
    t.oB = (t.intB ? (unsigned)fl_idx(g, x0, yB, 0) : safe) + (unsigned)ps * plane;
    {   // the thread's zero-bit word: row pair yA / 2 (yA is even), octet of the quad; clamped for rows / lanes outside the grid
        const int pairs = (int)(A.zb_plane / (unsigned)A.zb_lanes);
        const int zp = min(max(yA, 0) >> 1, pairs - 1), zo = min((x0 - 1) >> 3, A.zb_lanes - 1);
        t.zbv = A.zbits[b == 0 ? 0 : 1]; t.zoff = (unsigned)(zp * A.zb_lanes + zo); t.zsh = ((x0 - 1) >> 2 & 1) * 4;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { t.QA[i] = zero4; t.QB[i] = zero4; t.RA[i] = zero4; t.RB[i] = zero4; t.FA[i] = 0; }

    // prologue: own column of the plane behind ps (ring index 1) and of ps (index 2); flag bytes
    // of the first plane that is stored
    auto consume = [&]() -> const float* {                       // wait for the next plane in stream order
        sw_mbar_wait(&bars[s_con], ph_con);
        const float* s = uring + s_con * SLOTF;
        if (++s_con == NSLOT) { s_con = 0; ph_con ^= 1u; }
        return s;
    };
    if (has_behind) { const float* Ub = consume(); t.QA[1] = sw_lds4(Ub + t.offA); t.QB[1] = sw_lds4(Ub + t.offB); }
    const float* Up = consume();
    t.QA[2] = sw_lds4(Up + t.offA); t.QB[2] = sw_lds4(Up + t.offB);
    t.FA[3] = t.zbv[(unsigned)(ps + DIR + FL_HALO - 1) * A.zb_plane + t.zoff];
    __syncthreads();        // the slot of the plane behind ps is re-armed by the first step

    int p = ps, step = 0;
    int jc = has_behind ? 1 : 0;                                 // stream index of plane p
    // No proxy fence before the barrier: the slot the next TMA overwrites was only READ through
    // the generic proxy, and those reads are ordered before the issuing thread's TMA by the barrier.
#define SW_STEP(U, E)                                                                                       \
    {                                                                                                   \
        if (step == nsteps) break;                                                                      \
        if (issued < nplanes && issued - jc < NSLOT) {        /* a slot behind plane p is free */       \
            issue(first + DIR * issued);                                                                \
            ++issued;                                                                                   \
        }                                                                                               \
        t.oA += dplane; t.oB += dplane;                                                                 \
        const bool has_next = DIR > 0 ? (p + 1 <= hi) : (p - 1 >= lo);                                  \
        unsigned long long* bar_n = &bars[s_con];                                                       \
        const unsigned ph_n = ph_con;                                                                   \
        const float* Un = uring + s_con * SLOTF;                                                        \
        if (has_next) { if (++s_con == NSLOT) { s_con = 0; ph_con ^= 1u; } }                            \
        sw_step<BOXF, XSLOTF, MODE, DIR, U, E>(t, A, out, b, p, za, zb, has_next, Up, Un, bar_n,         \
                                               ph_n, vbuf);                                             \
        Up = Un; p += DIR; ++step; ++jc;                                                                \
        __syncthreads();                                                                                \
    }
    // A face segment streams away from its slab face: after four steps the two boundary planes are final
    // and go to the z-neighbour while the rest of the segment is still being swept.
    const int face = (sg.mode >> 1) - 1;
    // Four steps later the stores are long on their way: fence them and count this CTA in (the last one raises the
    // neighbour's flag) -- early in the segment, however long it is, so that the neighbour's next sweep never waits
    // for the end of this one (dbg 16: only when the segment is done, as before).
    bool signalled = false;
#define SW_HAND_OVER if (face >= 0 && !(A.dbg & 8)) {                                                                                  \
        if (step == 4) sw_push_face<GE::TY, GE::THREADS>(A, g, out, batch, sg.col - bx * A.nty, y0, xtile == 0 ? 0 : (FL_XOFF + own_lo) / 4,  \
                                                         xtile == A.nxt - 1 ? g.pitch / 4 : (FL_XOFF + own_hi + 1) / 4, face, tid);         \
        else if (step == 8 && !(A.dbg & 16)) { sw_signal_face(A, face, tid); signalled = true; } }
    if (((g.zlo + ps) & 1) == 0) { for (;;) { SW_HAND_OVER SW_STEP(0, 0) SW_STEP(1, 1) SW_STEP(2, 0) SW_STEP(3, 1) } }
    else                         { for (;;) { SW_HAND_OVER SW_STEP(0, 1) SW_STEP(1, 0) SW_STEP(2, 1) SW_STEP(3, 0) } }
#undef SW_STEP
#undef SW_HAND_OVER
    if (face >= 0 && !(A.dbg & 8) && !signalled) sw_signal_face(A, face, tid);
}

// MODE 0 = Jacobi, 1 = red-black (colour 0 = (x+y+z) even first).  A CTA works through a short list
// of segments (column = batched field x row tile, planes za..zb), each streamed in the direction the
// plan gave it, reversed when A.flip is set.  The host flips every other sweep: the planes a CTA wrote
// last are the ones it reads first in the next sweep, while they are still in L2.
// DIRS: 0 = every segment upwards, 1 = every segment downwards (single-GPU balanced plan: half the code,
// measurably fewer instruction-fetch stalls), 2 = per segment as described above.
template <int TXW, int NP, int CPS, int MODE, int DIRS>
__global__ void __launch_bounds__(32 * TXW * NP, CPS) k_sweep_stream(const __grid_constant__ StreamArgs A)
{
    using GE = SwGeom<TXW, NP, CPS>;
    constexpr int BOXF = GE::BOXF, RBOXF = GE::RBOXF, SLOTF = GE::SLOTF, XSLOTF = GE::XSLOTF;
    constexpr int NSLOT = GE::NSLOT;
    extern __shared__ __align__(128) float sw_mem[];
    __shared__ unsigned long long bars[NSLOT];
    float* const uring = sw_mem;                                     // NSLOT TMA slots (x boxes + rhs boxes)
    float* const vbuf = sw_mem + NSLOT * SLOTF;                      // 2 slots: plane after stage 0

    const Grid& g = A.g;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wx = warp % TXW, pr = warp / TXW;

    if (tid == 0) {
        for (int i = 0; i < NSLOT; ++i) sw_mbar_init(&bars[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Programmatic dependent launch (no-ops unless the launch carries the attribute): the next sweep of the
    // solve may be scheduled while this one drains, and this one touches no field data before the sweep it
    // depends on has completed and flushed.  Only the plan tables (written once, long ago) are read earlier.
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    __syncthreads();
    if (A.trace && tid == 0) { unsigned long long tt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt)); A.trace[4 * blockIdx.x] = tt; }

    // what is constant for the thread across segments
    SwThread t;
    t.lane = lane; t.sink = 0;
    const int qx = 32 * wx + lane;                                   // this thread's quad within the CTA's x extent
    const int rA = 1 + 2 * pr;
    t.offA = wx * BOXF + rA * SW_BOXW + 4 * (lane + 1); t.offB = t.offA + SW_BOXW;
    t.roffA = XSLOTF + wx * RBOXF + (rA - 1) * SW_RBOXW + 4 * lane; t.roffB = t.roffA + SW_RBOXW;
    // warp-edge lanes publish their outer cell into the neighbour column's halo, or copy the x ghosts
    t.pub_gl = lane == 0 && wx == 0;        t.pub_nl = lane == 0 && wx > 0;
    t.pub_gr = lane == 31 && wx == TXW - 1; t.pub_nr = lane == 31 && wx + 1 < TXW;
    const unsigned safe = (unsigned)fl_idx(g, 1, 1, 0);

    int s_iss = 0;                                                   // ring slot the next TMA plane goes to
    int s_con = 0; unsigned ph_con = 0;                              // slot / mbarrier parity of the next plane consumed

    for (int si = A.seg_first[blockIdx.x]; si < A.seg_first[blockIdx.x + 1]; ++si) {
        const SwSeg sg = A.segs[si];
        // face segments always stream away from their face (lower face: upwards), the others as planned,
        // reversed on odd sweeps
        const bool down = DIRS == 0 ? false : DIRS == 1 ? true
                        : (sg.mode >> 1) ? (sg.mode >> 1) == 2 : (((sg.mode ^ A.flip) & 1) != 0);
        if (down)
            sw_segment<TXW, NP, CPS, MODE, -1>(A, sg, t, uring, vbuf, bars, s_iss, s_con, ph_con, tid, pr, rA, qx, safe);
        else
            sw_segment<TXW, NP, CPS, MODE, 1>(A, sg, t, uring, vbuf, bars, s_iss, s_con, ph_con, tid, pr, rA, qx, safe);
        if (A.trace && tid == 0 && si == A.seg_first[blockIdx.x]) {
            unsigned long long tt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
            A.trace[4 * blockIdx.x + 1] = tt; A.trace[4 * blockIdx.x + 3] = (unsigned long long)(unsigned)sg.mode | ((unsigned long long)(sg.zb - sg.za + 1) << 8);
        }
    }
    if (A.trace && tid == 0) { unsigned long long tt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt)); A.trace[4 * blockIdx.x + 2] = tt; }
    if (A.nty < 0) A.dst[0][0] = __uint_as_float(t.sink);            // never taken: keeps t.sink observable
}

// ---- host side -------------------------------------------------------------------------------
typedef CUresult (*sw_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                 CUtensorMapFloatOOBfill);

struct SwPlan {                            // segment table of one (fields, row tiles) decomposition, on the device
    SwSeg* segs = nullptr; int* seg_first = nullptr; int nctas = 0;
};

#define SW_EDGE_DEFAULT 8
struct StreamState {
    sw_encode_fn encode = nullptr;
    std::map<std::pair<const float*, int>, CUtensorMap> maps;   // (field buffer, box rows * 1024 + box width) -> 3-D tensor map
    std::map<std::pair<int, int>, SwPlan> plans;                // (fields, owned rows per tile) -> segment table
    int zc_override = 0;
    bool zigzag = true, unbatch = false, overlap = true;
    int cfg = 0, dbg = 0;
    int face_mode = 2;                                          // FLUID_SW_FACE_CTA: 2 = edge segments first in the CTAs that would otherwise start behind them (plan part 6, one wave), 1 = separate edge CTAs (part 4), 0 = hand-over inside ordinary segments (part 5)
    int edge = SW_EDGE_DEFAULT;                                 // planes per edge segment at an interior slab face (>= 2)
    unsigned long long* trace = nullptr; const char* trace_path = nullptr; long trace_launch = 0, trace_at[2] = { -1, -1 };   // FLUID_SW_TRACE=path, FLUID_SW_TRACE_AT=i,j
    int segcost = SW_SEGCOST, last_w32 = 32;                    // FLUID_SW_SEGCOST, FLUID_SW_LASTW: the plan's charge per segment start (steps) and the cost of a plane of a ragged last row tile (1/32 of a plane elsewhere)
    int late_extra = -1;                                        // FLUID_SW_LATE: extra pipeline steps the plan charges a CTA that starts behind an edge CTA
    int plan_mode = 2, plan_min_fill = 85;
    // zero-bit volume (k_build_zbits): rebuilt whenever the flag volume was
    unsigned short* zbits[2] = { nullptr, nullptr }; int zb_pairs = 0, zb_lanes = 0; unsigned long long zb_version = ~0ull;
    int pdl = 1;                                                // FLUID_SW_PDL: programmatic dependent launch between the sweeps of a solve; 0 = off, 1 = on one GPU (measured: 22.78 -> 22.41 ms per 256^3 step), 2 = on z-slabs too                      // FLUID_SW_PLAN: 0 = auto, 1 = uniform ranges, 2 = balanced budget
};

inline StreamState* sw_state(fluid_handle* h)
{
    if (!h->tiled) {
        StreamState* s = new StreamState();
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            s->encode = (sw_encode_fn)fn;
        if (const char* e = getenv("FLUID_TILED_ZC")) s->zc_override = atoi(e);
        if (const char* e = getenv("FLUID_SW_ZIGZAG")) s->zigzag = atoi(e) != 0;
        if (const char* e = getenv("FLUID_SW_UNBATCH")) s->unbatch = atoi(e) != 0;
        if (const char* e = getenv("FLUID_SW_OVERLAP")) s->overlap = atoi(e) != 0;
        if (const char* e = getenv("FLUID_SW_CFG")) s->cfg = atoi(e);
        if (const char* e = getenv("FLUID_SW_DBG")) s->dbg = atoi(e);
        if (const char* e = getenv("FLUID_SW_EDGE")) { s->edge = atoi(e); if (s->edge < 2) s->edge = 2; }
        if (const char* e = getenv("FLUID_SW_FACE_CTA")) s->face_mode = atoi(e);
        if (const char* e = getenv("FLUID_SW_LATE")) s->late_extra = atoi(e);
        if (const char* e = getenv("FLUID_SW_SEGCOST")) { s->segcost = atoi(e); if (s->segcost < 1 || s->segcost > 15) s->segcost = SW_SEGCOST; }
        if (const char* e = getenv("FLUID_SW_LASTW")) { s->last_w32 = atoi(e); if (s->last_w32 < 8 || s->last_w32 > 32) s->last_w32 = 32; }
        if (const char* e = getenv("FLUID_SW_TRACE")) {
            s->trace_path = e;
            if (cudaMalloc(&s->trace, 4096 * 4 * sizeof(unsigned long long)) != cudaSuccess) s->trace = nullptr;
            if (const char* at = getenv("FLUID_SW_TRACE_AT")) sscanf(at, "%ld,%ld", &s->trace_at[0], &s->trace_at[1]);
        }
        if (const char* e = getenv("FLUID_SW_PLAN")) s->plan_mode = atoi(e);
        if (const char* e = getenv("FLUID_SW_PLAN_FILL")) s->plan_min_fill = atoi(e);
        if (const char* e = getenv("FLUID_SW_PDL")) s->pdl = atoi(e);
        h->tiled = s;
    }
    return (StreamState*)h->tiled;
}

int fl_pair_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid, float a, float c,
                  unsigned x_from_rhs, int* done);

inline void fl_tiled_release(fluid_handle* h)
{
    StreamState* s = (StreamState*)h->tiled;
    if (s) for (auto& kv : s->plans) { cudaFree(kv.second.segs); cudaFree(kv.second.seg_first); }
    if (s) { cudaFree(s->zbits[0]); cudaFree(s->zbits[1]); cudaFree(s->trace); }
    delete s;
    h->tiled = nullptr;
}

// One bit per cell: "setBounds zeroes this cell" for scalars ([0]: solid) and for velocity components
// ([1]: solid or touching a solid), simulation.cpp:219-245.  Layout: one 16-bit word per (local plane
// 1-FL_HALO .. Dl+FL_HALO, row pair, octet): bits 0-7 = cells 1 + 8*octet .. + 7 of row 2*pair, bits 8-15 = the
// same cells of row 2*pair + 1.  2 x 2 MB at 256^3: the sweeps read them from L2 instead of one flag byte per
// cell from HBM (19.8 MB per field-sweep at 256^3).
__global__ void __launch_bounds__(256) k_build_zbits(Grid g, const unsigned char* __restrict__ flags,
                                                     unsigned short* __restrict__ z0, unsigned short* __restrict__ z1,
                                                     int pairs, int lanes)
{
    const long long n = (long long)(g.Dl + 2 * FL_HALO) * pairs * lanes;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int oct = (int)(t % lanes);
        const long long r = t / lanes;
        const int pr = (int)(r % pairs), zl = (int)(r / pairs) - (FL_HALO - 1);
        unsigned s0 = 0u, s1 = 0u;
        // the flag volume is non-zero exactly where a cell of the tunnel's interior is held with its z-neighbours
        if (zl >= 2 - FL_HALO && zl <= g.Dl + FL_HALO - 1)
            for (int k = 0; k < 2; ++k) {
                const int y = 2 * pr + k;
                if (y < 1 || y > g.H) continue;
                for (int i = 0; i < 8; ++i) {
                    const int x = 1 + 8 * oct + i;
                    if (x > g.W) break;
                    const unsigned f = flags[fl_idx(g, x, y, zl)];
                    if (f & FL_SOLID) s0 |= 1u << (8 * k + i);
                    if (f & (FL_SOLID | FL_NOSLIP)) s1 |= 1u << (8 * k + i);
                }
            }
        z0[t] = (unsigned short)s0; z1[t] = (unsigned short)s1;
    }
}

inline int sw_get_zbits(fluid_handle* h)
{
    StreamState* s = sw_state(h);
    const Grid& g = h->g;
    if (!s->zbits[0]) {
        s->zb_pairs = (g.rows + 1) / 2;
        s->zb_lanes = ((g.W + 7) / 8 + 31) / 32 * 32;
        const size_t n = (size_t)(g.Dl + 2 * FL_HALO) * s->zb_pairs * s->zb_lanes * sizeof(unsigned short);
        FL_CUDA(h, cudaMalloc(&s->zbits[0], n));
        FL_CUDA(h, cudaMalloc(&s->zbits[1], n));
    }
    if (s->zb_version != h->flags_version) {
        const long long n = (long long)(g.Dl + 2 * FL_HALO) * s->zb_pairs * s->zb_lanes;
        k_build_zbits<<<fl_grid_1d(n), 256, 0, h->stream>>>(g, h->cflags, s->zbits[0], s->zbits[1], s->zb_pairs, s->zb_lanes);
        FL_CHECK_LAUNCH(h);
        s->zb_version = h->flags_version;
    }
    return 0;
}

inline int sw_get_map(fluid_handle* h, const float* buf, int box_w, int box_rows, CUtensorMap* out)
{
    StreamState* s = sw_state(h);
    auto key = std::make_pair(buf, box_rows * 1024 + box_w);
    auto it = s->maps.find(key);
    if (it == s->maps.end()) {
        const Grid& g = h->g;
        CUtensorMap m;
        cuuint64_t dims[3] = { (cuuint64_t)g.pitch, (cuuint64_t)g.rows, (cuuint64_t)(g.Dl + 2 + 2 * FL_ZPAD) };   // planes -FL_ZPAD .. Dl+1+FL_ZPAD
        cuuint64_t strides[2] = { (cuuint64_t)g.pitch * 4, (cuuint64_t)g.plane * 4 };
        cuuint32_t box[3] = { (cuuint32_t)box_w, (cuuint32_t)box_rows, 1 };
        cuuint32_t estr[3] = { 1, 1, 1 };
        CUresult r = s->encode(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)(buf - FL_ZPAD * g.plane), dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fl_fail(h, FLUID_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
        it = s->maps.emplace(key, m).first;
    }
    *out = it->second;
    return 0;
}

inline std::pair<int, int> sw_plan_key(int nb, int ty, int cps, int part, int segcost) { return std::make_pair(nb, ((ty * 16 + cps) * 8 + part) * 16 + segcost); }

// the plan of sweep_plan.h for this handle, built once per (fields, tile, part) and kept on the device
inline int sw_get_plan(fluid_handle* h, int nb, int ty, int cps, int part, SwPlan* out, int segcost = SW_SEGCOST)
{
    StreamState* s = sw_state(h);
    auto key = sw_plan_key(nb, ty, cps, part, segcost);
    auto it = s->plans.find(key);
    if (it == s->plans.end()) {
        SwPlanParams p;
        p.Dl = h->g.Dl; p.H = h->g.H; p.nb = nb; p.ty = ty; p.cps = cps; p.part = part; p.segcost = segcost;
        p.sm_count = h->sm_count; p.lo_face = h->rank > 0; p.hi_face = h->rank < h->nranks - 1;
        p.edge = s->edge; p.late_extra = s->late_extra; p.zc_override = s->zc_override;
        p.plan_mode = s->plan_mode; p.plan_min_fill = s->plan_min_fill;
        p.last_w32 = s->last_w32;
        std::vector<SwSeg> segs; std::vector<int> first;
        sw_build_plan(p, segs, first);
        SwPlan pl;
        pl.nctas = (int)first.size() - 1;
        FL_CUDA(h, cudaMalloc(&pl.segs, segs.size() * sizeof(SwSeg)));
        FL_CUDA(h, cudaMalloc(&pl.seg_first, first.size() * sizeof(int)));
        FL_CUDA(h, cudaMemcpyAsync(pl.segs, segs.data(), segs.size() * sizeof(SwSeg), cudaMemcpyHostToDevice, h->stream));
        FL_CUDA(h, cudaMemcpyAsync(pl.seg_first, first.data(), first.size() * sizeof(int), cudaMemcpyHostToDevice, h->stream));
        FL_CUDA(h, cudaStreamSynchronize(h->stream));    // the host vectors go out of scope
        it = s->plans.emplace(key, pl).first;
    }
    *out = it->second;
    return 0;
}

template <int TXW, int NP, int CPS>
inline int sw_run(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid, float a, float c,
                  bool jacobi, unsigned x_from_rhs, int nsweeps, int first_sweep = 0)
{
    using GE = SwGeom<TXW, NP, CPS>;
    const Grid& g = h->g;
    StreamState* s = sw_state(h);
    // one streaming direction per launch unless segments carry their own (z-slabs, uniform-range plan)
    const bool mixed = h->nranks > 1 || s->plan_mode != 2;
    typedef void (*sweep_kernel)(const StreamArgs);
    const sweep_kernel kerns[3] = {
        jacobi ? k_sweep_stream<TXW, NP, CPS, 0, 0> : k_sweep_stream<TXW, NP, CPS, 1, 0>,
        jacobi ? k_sweep_stream<TXW, NP, CPS, 0, 1> : k_sweep_stream<TXW, NP, CPS, 1, 1>,
        jacobi ? k_sweep_stream<TXW, NP, CPS, 0, 2> : k_sweep_stream<TXW, NP, CPS, 1, 2> };
    for (int i = 0; i < 3; ++i)
        FL_CUDA(h, cudaFuncSetAttribute(kerns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GE::SMEM));
    // z-slabs.  The planes next to the slab faces are swept by CTAs of their own, ahead of the planes in
    // between, so that the halo hand-over overlaps the rest of the sweep.
    //  * fused (default once the neighbours' buffers are IPC-mapped): one launch per sweep; the edge CTAs
    //    store their boundary planes into the neighbour's halo over NVLink and raise its arrival flag, the
    //    next sweep's launch waits on this rank's flags with a stream memory operation.  No second
    //    stream, no communication kernel.
    //  * NCCL: the edge CTAs run on the high-priority halo stream, ncclSend/ncclRecv follows them there.
    // Slabs too thin for separate edge CTAs sweep whole columns and exchange afterwards.
    const bool overlap = h->nranks > 1 && s->overlap && s->zc_override == 0 && g.Dl >= 4 * s->edge;
    const bool push = h->nranks > 1 && fl_push_usable(h);
    const bool fused = overlap && push;
    // x tiles: a grid wider than the CTA's 128 * TXW cells is cut into tiles of equal width, each with one halo
    // quad beside every interior edge (x is read 1 + 2/62 times instead of once; the alternative -- a CTA as wide as
    // the grid with two owned rows -- read it three times).  The plan sees (field, x tile) pairs as its "fields".
    const int nq = g.W / 4, qmax = 32 * TXW;
    int nxt = 1, xq = nq;
    if (nq > qmax) { nxt = (nq - 2 + (qmax - 2) - 1) / (qmax - 2); xq = (nq + nxt - 1) / nxt; }
    const int nbx = nb * nxt;
    SwPlan plan, plan_edge;
    const int plan_part = fused ? (s->face_mode == 2 ? 6 : s->face_mode == 1 ? 4 : 5) : overlap ? 2 : 0;
    if (int e = sw_get_plan(h, nbx, GE::TY, CPS, plan_part, &plan, s->segcost)) return e;
    if (overlap && !fused) if (int e = sw_get_plan(h, nbx, GE::TY, CPS, 1, &plan_edge)) return e;
    StreamArgs A;
    A.g = g; A.a = a; A.c_recip = 1.0f / c;                              // simulation.cpp:257
    if (int e = sw_get_zbits(h)) return e;
    A.zbits[0] = s->zbits[0]; A.zbits[1] = s->zbits[1]; A.zb_lanes = s->zb_lanes; A.zb_plane = (unsigned)(s->zb_pairs * s->zb_lanes);
    A.nty = (g.H + GE::TY - 1) / GE::TY;
    A.nxt = nxt; A.xt_cells = 4 * xq;
    A.segs = plan.segs; A.seg_first = plan.seg_first;
    A.dbg = s->dbg; A.trace = nullptr;

    float* cur[4]; float* alt[4];
    if (nb > 4) return fl_fail(h, FLUID_E_ARG, "streaming solve: %d batched fields", nb);
    if (nb == 4 && !h->scratch[3]) {                                      // fourth ping-pong partner, first use
        const size_t fb = (size_t)g.cells * sizeof(float);
        FL_CUDA(h, cudaMalloc(&h->scratch[3], fb));
        FL_CUDA(h, cudaMemsetAsync(h->scratch[3], 0, fb, h->stream));
        h->scratch[3] += FL_ZPAD * g.plane;
    }
    for (int i = 0; i < nb; ++i) {
        cur[i] = h->field[fid[i]]; alt[i] = h->scratch[i];
        A.b[i] = b[i];
        if (int e = sw_get_map(h, h->field[pid[i]], SW_RBOXW, GE::BR - 2, &A.rmap[i])) return e;
    }
    dim3 ge((2 * (g.W + 2) + 2 * g.H + 127) / 128, g.Dl + 2);
    for (int i = 0; i < nb; ++i) {                                        // edge/corner ghosts follow
        if ((x_from_rhs >> i) & 1u) {                                     // field[fid] is a stale buffer: both partners take them from the guess
            k_copy_edges<<<ge, 128, 0, h->stream>>>(g, h->field[pid[i]], cur[i]);
            FL_CHECK_LAUNCH(h);
            k_copy_edges<<<ge, 128, 0, h->stream>>>(g, h->field[pid[i]], alt[i]);
        } else
            k_copy_edges<<<ge, 128, 0, h->stream>>>(g, cur[i], alt[i]);
        FL_CHECK_LAUNCH(h);
    }
    // Single-field solves re-read the same right-hand side acc times: optionally pin it in the
    // persisting part of L2 (off by default, see fluid_create).
    const size_t fbytes = (size_t)g.cells * sizeof(float);
    const bool pin = nb == 1 && h->l2_persist_max > 0 && nsweeps > 1;
    if (pin) {
        cudaStreamAttrValue av = {};
        av.accessPolicyWindow.base_ptr = (void*)h->field[pid[0]];
        av.accessPolicyWindow.num_bytes = fbytes < h->l2_window_max ? fbytes : h->l2_window_max;
        const double r = (double)h->l2_persist_max * 0.9 / (double)av.accessPolicyWindow.num_bytes;
        av.accessPolicyWindow.hitRatio = r > 1.0 ? 1.0f : (float)r;
        av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        FL_CUDA(h, cudaStreamSetAttribute(h->stream, cudaStreamAttributeAccessPolicyWindow, &av));
    }
    for (int k = 0; k < nsweeps; ++k) {                                   // simulation.cpp:258
        // x_from_rhs: the first sweep reads its guess from the right-hand-side buffer (same values: the caller
        // swapped the two fields instead of copying, simulation.cpp:108-110) and writes field[fid]
        const unsigned from_rhs = k == 0 ? x_from_rhs : 0u;                // per batched field
        for (int i = 0; i < nb; ++i) {
            const bool fr = (from_rhs >> i) & 1u;
            if (int e = sw_get_map(h, fr ? h->field[pid[i]] : cur[i], SW_BOXW, GE::BR, &A.map[i])) return e;
            A.dst[i] = fr ? cur[i] : alt[i];
        }
        A.flip = (s->zigzag && ((k + first_sweep) & 1)) ? 1 : 0;         // odd sweeps stream the other way (L2 reuse)
        const sweep_kernel kern = kerns[mixed ? 2 : A.flip];
        const bool tracing = s->trace && s->trace_path && plan.nctas <= 4096 && (s->trace_launch == s->trace_at[0] || s->trace_launch == s->trace_at[1]);
        A.trace = tracing ? s->trace : nullptr;
        ++s->trace_launch;
        const unsigned seq = push ? ++h->halo_seq : 0u;
        if (fused) {
            if (int e = fl_push_targets(h, nb, A.dst, A.peer_dst, A.peer_flag, A.peer_zdst)) return e;   // the buffers this sweep writes
            A.counter = h->d_sync + 2; A.n_edge = nbx * A.nty; A.seq = seq;
            A.my_flag = h->d_sync; A.err_word = h->d_sync + 16;
        } else if (overlap) {
            // ev_a = "the in-between planes of the previous sweep (or whatever preceded the solve) are done"
            if (k == 0) FL_CUDA(h, cudaEventRecord(h->ev_a, h->stream));
            FL_CUDA(h, cudaStreamWaitEvent(h->halo_stream, h->ev_a, 0));
            StreamArgs E = A; E.segs = plan_edge.segs; E.seg_first = plan_edge.seg_first;
            kern<<<plan_edge.nctas, GE::THREADS, GE::SMEM, h->halo_stream>>>(E);
            FL_CHECK_LAUNCH(h);
        }
        if ((s->pdl == 1 && h->nranks == 1) || s->pdl == 2) {
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
        if (tracing) {                                                     // diagnostic: per-CTA timeline of this launch
            std::vector<unsigned long long> tr((size_t)plan.nctas * 4);
            FL_CUDA(h, cudaStreamSynchronize(h->stream));
            FL_CUDA(h, cudaMemcpy(tr.data(), s->trace, tr.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
            char path[512]; snprintf(path, sizeof path, "%s.r%d.l%ld.txt", s->trace_path, h->rank, s->trace_launch - 1);
            if (FILE* f = fopen(path, "w")) {
                unsigned long long t0 = ~0ull;
                for (int k = 0; k < plan.nctas; ++k) if (tr[4 * k] < t0) t0 = tr[4 * k];
                fprintf(f, "# rank %d launch %ld fields %d ctas %d: cta start_ns first_segment_done_ns end_ns first_mode first_planes\n", h->rank, s->trace_launch - 1, nb, plan.nctas);
                for (int k = 0; k < plan.nctas; ++k)
                    fprintf(f, "%d %llu %llu %llu %llu %llu\n", k, tr[4 * k] - t0, tr[4 * k + 1] - t0, tr[4 * k + 2] - t0, tr[4 * k + 3] & 0xffull, tr[4 * k + 3] >> 8);
                fclose(f);
            }
        }
        for (int i = 0; i < nb; ++i) if (!((from_rhs >> i) & 1u)) { float* t = cur[i]; cur[i] = alt[i]; alt[i] = t; }
        if (fused) {
            // the next sweep's edge CTAs wait for the neighbours' halos themselves (sw_await_face);
            // whatever follows the solve waits once, below
        } else if (overlap) {
            if (int e = fl_halo_exchange_on(h, nb, cur, h->halo_stream)) return e;   // two planes each way
            FL_CUDA(h, cudaEventRecord(h->ev_b, h->halo_stream));
            FL_CUDA(h, cudaEventRecord(h->ev_a, h->stream));
            FL_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_b, 0));                  // next sweep sees the halos
        } else if (push) {
            if (int e = fl_halo_push(h, nb, cur, h->stream, seq)) return e;
            if (int e = fl_halo_wait(h, h->stream, seq)) return e;
        } else if (h->nranks > 1) {
            if (int e = fl_halo_exchange(h, nb, cur)) return e;
        }
    }
    if (fused && nsweeps > 0) if (int e = fl_halo_wait(h, h->stream, h->halo_seq)) return e;
    if (pin) {
        cudaStreamAttrValue av = {};
        av.accessPolicyWindow.num_bytes = 0;
        FL_CUDA(h, cudaStreamSetAttribute(h->stream, cudaStreamAttributeAccessPolicyWindow, &av));
    }
    for (int i = 0; i < nb; ++i)
        if (cur[i] != h->field[fid[i]]) { h->scratch[i] = h->field[fid[i]]; h->field[fid[i]] = cur[i]; }
    return 0;
}

// does the streaming kernel take this handle's solves?  (shapes it declines go to the simple kernels)
inline bool fl_tiled_eligible(fluid_handle* h)
{
    const Grid& g = h->g;
    const unsigned solver = h->flags & FLUID_SOLVER_MASK;
    if (h->flags & FLUID_NO_TILED) return false;
    if (solver != FLUID_SOLVER_RB && solver != FLUID_SOLVER_JACOBI) return false;
    if ((g.W & 3) || g.W > 8192) return false;            // quads must not straddle the x = W+1 ghost
    if (!(h->flags & FLUID_TILED_ALWAYS) && (g.W < 64 || g.H < 8 || g.Dl < 4)) return false;
    return true;
}

// step() may swap v <-> v_prev instead of copying (simulation.cpp:108-110) when the solve that follows is an
// out-of-place streaming solve on one GPU: its first sweep then reads the guess from the right-hand side
inline bool fl_swap_instead_of_copy(fluid_handle* h)
{
    static const bool off = getenv("FLUID_NO_SWAP") != nullptr;
    static const bool slab_off = getenv("FLUID_NO_SLAB_SWAP") != nullptr;
    if (off || h->acc <= 0 || !fl_tiled_eligible(h)) return false;
    // z-slabs: every rank swaps the same way, and the peers' mappings are looked up by allocation, so the
    // hand-over follows the buffers; only on the path where the sweeps push their halos themselves
    StreamState* s = sw_state(h);
    bool mine = s->encode != nullptr && !s->unbatch;
    if (h->nranks > 1) {
        if (h->slab_swap_ok >= 0) return h->slab_swap_ok != 0;
        if (slab_off || !fl_push_usable(h)) mine = false;
        // advect() gathers from every rank's *_prev buffers, which after a swap are the allocations that held v:
        // both sets must be mapped from everybody (hosts that map only the original gather sources keep the copies)
        for (int r = 0; r < h->nranks && mine; ++r)
            for (int k = 0; k < 3 && r != h->rank; ++k)
                if (!fl_peer_mapped(h, r, h->field[FLUID_VX + k]) || !fl_peer_mapped(h, r, h->field[FLUID_VX_PREV + k])) mine = false;
        // the ranks must not differ in this (a rank that copies while its neighbour swaps reads the wrong buffers):
        // asked once per handle, at its first step -- a collective like the step itself
        bool all = false;
        if (fl_all_agree(h, mine, &all) != 0) all = false;
        h->slab_swap_ok = all ? 1 : 0;
        return all;
    }
    return mine;
}

// a z-slab runs the density diffusion (simulation.cpp:135) as the fourth field of the velocity diffusion when the
// solve is certain to go through the streaming kernel with the peer-store hand-over (four mapped partners)
bool fl_slab_batch4(fluid_handle* h)
{
    if (h->nranks == 1 || h->acc <= 0 || !fl_tiled_eligible(h) || !fl_push_usable(h)) return false;
    StreamState* s = sw_state(h);
    return s->encode != nullptr && !s->unbatch && h->scratch[3] != nullptr;
}

// linearSolver through the streaming kernel; *handled = false leaves the solve to the simple kernels
inline int fl_tiled_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                          float a, float c, bool* handled)
{
    *handled = false;
    const Grid& g = h->g;
    const unsigned solver = h->flags & FLUID_SOLVER_MASK;
    const unsigned x_from_rhs = h->x_from_rhs;
    h->x_from_rhs = 0;
    if (!fl_tiled_eligible(h)) return x_from_rhs ? fl_fail(h, FLUID_E_STATE, "internal: swapped fields reached a non-streaming solve") : 0;
    StreamState* s = sw_state(h);
    if (!s->encode) return fl_fail(h, FLUID_E_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
    const bool jac = solver == FLUID_SOLVER_JACOBI;
    // batched fields either go out as one launch per sweep, or (FLUID_SW_UNBATCH=1) one field after
    // the other so that a sweep's footprint is a third as large and more of it is still in L2
    const int ngroups = (s->unbatch && nb > 1) ? nb : 1, per = nb / ngroups;
    for (int gi = 0; gi < ngroups; ++gi) {
        const int* bb = b + gi * per; const int* ff = fid + gi * per; const int* pp = pid + gi * per;
        int e = 0, done = 0;
        unsigned xfr = x_from_rhs;
        // two sweeps per HBM pass where the temporally blocked kernel applies (sweep_pair.cuh); an odd
        // sweep count ends with one ordinary sweep
        if (!jac && per <= 4) { e = fl_pair_solve(h, per, bb, ff, pp, a, c, xfr, &done); if (e) return e; if (done) xfr = 0u; }
        const int rest = h->acc - done;
        if (rest <= 0) continue;
        // W <= 128: one warp column, 30 owned rows; <= 256: two, 14 rows; <= 512: four, 6 rows; wider grids: the
        // two-column CTA in x tiles (measured at 1024 x 512 x 64: 112 us per field-sweep, 0.55 of the roofline, against
        // 194 us / 0.32 for round 1's eight-column CTA with two owned rows; at W = 512 the tiles fill only 45 of 64
        // lanes and lose to the four-column CTA, 0.54 against 0.69 -- FLUID_SW_CFG=5 forces them there for A/B)
        if (g.W <= 128)      e = sw_run<1, 16, 1>(h, per, bb, ff, pp, a, c, jac, xfr, rest, done);
        else if (s->cfg == 1 && g.W <= 256) e = sw_run<2, 4, 2>(h, per, bb, ff, pp, a, c, jac, xfr, rest, done);
        else if (s->cfg == 2 && g.W <= 256) e = sw_run<2, 5, 2>(h, per, bb, ff, pp, a, c, jac, xfr, rest, done);
        else if (s->cfg == 3 && g.W <= 256) e = sw_run<2, 6, 1>(h, per, bb, ff, pp, a, c, jac, xfr, rest, done);
        else if (g.W > 256 && g.W <= 512 && s->cfg != 5) e = sw_run<4, 4, 1>(h, per, bb, ff, pp, a, c, jac, xfr, rest, done);
        else                  e = sw_run<2, 8, 1>(h, per, bb, ff, pp, a, c, jac, xfr, rest, done);
        if (e) return e;
    }
    *handled = true;
    return 0;
}
