// sweep_plan.h -- which CTA of a streaming sweep takes which planes of which column: pure host code (no CUDA), so
// that tests/test_sweep_plan.py can check its properties for any shape on the CPU (tests/host_emul/plan_check.cpp).
#pragma once
#include <vector>

// column (batched field * row tiles + row tile), planes za..zb; mode bit 0 = streamed downwards,
// bits 1-2 = 1 + slab face (0 lower, 1 upper) whose two boundary planes this segment hands to the z-neighbour
struct SwSeg { int col, za, zb, mode; };

// Work decomposition.  The unit of work is one plane of one column (column = batched field x row
// tile); the ncol * Dl units, in (column, z) order, are cut into one contiguous range per SM, so
// every SM streams the same number of planes (+-1) whatever the tile count.  A range that crosses a
// column boundary becomes two segments; cuts closer than SW_MINSEG planes to a column boundary are
// snapped onto it, because every segment pays 4 extra pipeline steps.
#define SW_MINSEG 5
#define SW_SEGCOST 4
// part 0 = all planes 1..Dl.  A z-slab whose halo exchange is overlapped with the sweep splits them:
// part 1 = the `edge` planes next to each interior slab face (one CTA per column and face; they run
// first, on the halo stream, and the exchange follows them), part 2 = the planes in between on the SMs
// the edge CTAs leave free (NCCL's copy kernel needs SMs of its own once they are done), part 3 = the
// planes in between on all SMs (halos handed over by peer stores: nothing else needs an SM), where the
// CTAs that only start once an edge CTA has retired get that much less work; part 4 = parts 1 and 3 in
// one grid, edge CTAs first, for the sweep that hands its halos over itself; part 5 = all planes as in
// part 0, the segments that touch a slab face doing the hand-over on their way (FLUID_SW_FACE_CTA=0; slower with
// two faces: 4 GPUs 24.9 vs 24.1 ms/step); part 6 = part 4 in ONE wave: the edge segment is the first segment of
// the CTA that part 4 would start behind it, so no CTA retires and none is dispatched mid-sweep (the default:
// 22.6 vs 22.8 ms on 2 GPUs, 23.2 vs 23.4 on 4).
struct SwPlanParams {
    int Dl, H;                             // planes of the slab, rows of the grid
    int nb, ty, cps, part, segcost;        // columns per row tile (batched fields x x tiles), owned rows per tile, CTAs per SM
    int sm_count;
    bool lo_face, hi_face;                 // interior slab faces (a neighbour rank below / above)
    int edge, late_extra, zc_override, plan_mode, plan_min_fill;   // StreamState's knobs
    int last_w32 = 32;                     // cost of a plane of a ragged last row tile in 1/32 of a plane elsewhere (32: no difference)
};

// CTA k works through segs[first[k] .. first[k+1]).  Returns the part that was built (a part that cannot be cut
// for this shape falls back to part 4).
inline int sw_build_plan(const SwPlanParams& p, std::vector<SwSeg>& segs, std::vector<int>& first)
{
    const bool lo_face = p.lo_face, hi_face = p.hi_face;
    const int nb = p.nb, ty = p.ty, cps = p.cps, part = p.part, segcost = p.segcost;
    segs.clear(); first.clear();
    // planes zoff+1 .. zoff+Dl of every column are distributed (part 0: zoff = 0, Dl = p.Dl)
    const bool split = (part >= 2 && part <= 4) || part == 6;   // edge planes have segments of their own
    const int zoff = (split && lo_face) ? p.edge : 0;
    const int Dl = split ? p.Dl - zoff - (hi_face ? p.edge : 0) : p.Dl;
    const int nty = (p.H + ty - 1) / ty, ncol = nb * nty;
    const long long T = (long long)ncol * Dl;
    std::vector<SwSeg> edge_segs;                       // part 6: kept aside, merged below
    if (part == 1 || part == 4 || part == 6) {
        // part 4: the edge CTAs come first in the grid (they are dispatched first) and hand their
        // boundary planes to the z-neighbour themselves; they stream away from the face
        const int lo_mode = part != 1 ? (1 << 1) : 0, hi_mode = part != 1 ? (2 << 1) | 1 : 0;
        for (int c = 0; c < ncol; ++c) {
            if (part == 6) {
                if (lo_face) edge_segs.push_back(SwSeg{ c, 1, p.edge, lo_mode });
                if (hi_face) edge_segs.push_back(SwSeg{ c, p.Dl - p.edge + 1, p.Dl, hi_mode });
                continue;
            }
            if (lo_face) { first.push_back((int)segs.size()); segs.push_back(SwSeg{ c, 1, p.edge, lo_mode }); }
            if (hi_face) { first.push_back((int)segs.size()); segs.push_back(SwSeg{ c, p.Dl - p.edge + 1, p.Dl, hi_mode }); }
        }
    }
    if (part == 1) {
    } else
    if (p.zc_override > 0) {                        // tests: fixed chunk height, one CTA per segment
        const int zc = p.zc_override;
        for (int c = 0; c < ncol; ++c)
            for (int za = 1; za <= Dl; za += zc) {
                first.push_back((int)segs.size());
                segs.push_back(SwSeg{ c, zoff + za, zoff + (za + zc - 1 < Dl ? za + zc - 1 : Dl), 0 });
            }
    } else {
        long long nc = (long long)p.sm_count * cps;
        // overlapped exchange: the edge CTAs (part 1) run at the same time, and NCCL's copy kernel needs
        // SMs of its own once they are done -- leave them room instead of queueing behind a full wave
        if (part == 2) nc -= (long long)ncol * ((lo_face ? 1 : 0) + (hi_face ? 1 : 0));
        if (nc < p.sm_count / 2) nc = p.sm_count / 2;
        if (nc > T / 8) nc = T / 8;                  // tiny grids: at least 8 planes per CTA
        if (nc < 1) nc = 1;
        // parts 3 / 4: the edge CTAs are dispatched first, in waves of nc; an in-between CTA starts when a
        // slot frees up, so the last n_edge % nc of them start one edge segment (+ its pipeline fill) later
        // than the others and get that much less work
        const long long n_edge = (long long)ncol * ((lo_face ? 1 : 0) + (hi_face ? 1 : 0));
        const long long n_prompt = (part == 3 || part == 4 || part == 6) ? nc - n_edge % nc : nc;
        // measured (4 GPUs, 256^3 per rank): a rank with two faces gains 2.8 % when a late CTA is charged 4 steps more,
        // a rank with one face loses as much (2 GPUs) -- FLUID_SW_LATE overrides
        const long long late = p.edge + segcost + 2 + (p.late_extra >= 0 ? p.late_extra : (lo_face && hi_face) ? 4 : 0);
        // Uniform plan: every column is cut into the same S ranges and one CTA streams range k of a
        // column, upwards for even k, downwards for odd k.  Neighbouring row tiles are then on the same
        // plane at the same time and the ranges of a column reach their shared boundary together, so
        // halo rows and halo planes are L2 hits instead of second HBM reads.  It leaves nc - S*ncol
        // SMs idle, which the HBM-bound kernel tolerates up to a point (plan_min_fill, per cent).
        const long long S = nc / ncol;
        const bool uniform = part < 3 && (p.plan_mode == 1 ||
            (p.plan_mode == 0 && S >= 1 && S * ncol * 100 >= nc * p.plan_min_fill && Dl / (S > 0 ? S : 1) >= 3 * SW_MINSEG));
        if (uniform && S >= 1 && Dl / S >= 1) {
            for (long long k = 0; k < S; ++k)
                for (int c = 0; c < ncol; ++c) {
                    const int z0 = (int)(k * Dl / S), z1 = (int)((k + 1) * Dl / S);
                    first.push_back((int)segs.size());
                    segs.push_back(SwSeg{ c, zoff + z0 + 1, zoff + z1, (int)(k & 1) });
                }
        } else {
        // every CTA gets the same budget of pipeline steps: planes + SW_SEGCOST per segment it
        // starts.  The smallest budget with which nc CTAs cover all columns is found by bisection.
        // Accounting in 1/32 of a pipeline step: a plane of the last row tile of a field costs p.last_w32 / 32 of a
        // plane elsewhere when that tile owns fewer rows than the others (measured 0.92 at 4 of 14 rows: fewer stores).
        const bool ragged = p.last_w32 != 32 && p.H % ty != 0;
        auto weight = [&](int c) -> long long { return (ragged && c % nty == nty - 1) ? p.last_w32 : 32; };
        auto build = [&](long long budget, bool emit) -> bool {
            int c = 0, z = 0;
            for (long long k = 0; k < nc; ++k) {
                if (emit) first.push_back((int)segs.size());
                long long left = 32 * (budget - (k >= n_prompt ? late : 0));
                while (c < ncol) {
                    const int rem = Dl - z;
                    const long long w = weight(c), room = (left - 32 * segcost) / w;
                    long long take = room < rem ? room : rem;
                    if (rem - take > 0 && rem - take < SW_MINSEG) take = rem;       // no tiny leftovers
                    if (take < SW_MINSEG && take < rem) break;                      // no tiny starts
                    if (take <= 0) break;
                    if (emit) segs.push_back(SwSeg{ c, zoff + z + 1, zoff + z + (int)take, 0 });
                    left -= take * w + 32 * segcost; z += (int)take;
                    if (z == Dl) { ++c; z = 0; }
                }
            }
            return c == ncol;
        };
        long long lo_b = ragged ? 0 : T / nc, hi_b = T / nc + 3 * segcost + 2 * SW_MINSEG + Dl;
        while (!build(hi_b, false)) hi_b *= 2;
        while (lo_b + 1 < hi_b) {
            const long long mid = (lo_b + hi_b) / 2;
            if (build(mid, false)) hi_b = mid; else lo_b = mid;
        }
        build(hi_b, true);
        }
    }
    first.push_back((int)segs.size());
    if (part == 6) {
        // One wave: the CTA that part 4 starts behind edge CTA k sweeps that edge segment itself, first, and goes
        // on with its (shorter) share of the planes in between -- no CTA retires and none is dispatched mid-sweep.
        const long long nc_built = (long long)first.size() - 1;
        const long long n_e = (long long)edge_segs.size();
        if (p.zc_override > 0 || n_e > nc_built) { SwPlanParams q = p; q.part = 4; return sw_build_plan(q, segs, first); }
        std::vector<SwSeg> s2; std::vector<int> f2;
        for (long long k = 0; k < nc_built; ++k) {
            f2.push_back((int)s2.size());
            const long long e = k - (nc_built - n_e);
            if (e >= 0) s2.push_back(edge_segs[(size_t)e]);
            s2.insert(s2.end(), segs.begin() + first[(size_t)k], segs.begin() + first[(size_t)k + 1]);
        }
        f2.push_back((int)s2.size());
        segs.swap(s2); first.swap(f2);
    }
    if (part == 5) {
        // Hand-over inside ordinary segments: the first segment of a column (za = 1) takes the lower face,
        // the last (zb = Dl) the upper face; a segment that spans a whole column with two faces is cut
        // in the middle.  Face segments go first in their CTA, so the flags rise early in the sweep.
        std::vector<SwSeg> s2; std::vector<int> f2;
        bool ok = true;
        for (size_t k = 0; k + 1 < first.size(); ++k) {
            std::vector<SwSeg> mine, faces, rest;
            for (int i = first[k]; i < first[k + 1]; ++i) {
                SwSeg sg = segs[i];
                const bool lo = lo_face && sg.za == 1, hi = hi_face && sg.zb == p.Dl;
                if (lo && hi) {
                    const int m = (sg.za + sg.zb) / 2;
                    mine.push_back(SwSeg{ sg.col, sg.za, m, 1 << 1 });
                    mine.push_back(SwSeg{ sg.col, m + 1, sg.zb, (2 << 1) | 1 });
                } else {
                    if (lo) sg.mode = 1 << 1;
                    if (hi) sg.mode = (2 << 1) | 1;
                    mine.push_back(sg);
                }
            }
            for (const SwSeg& sg : mine) {
                if ((sg.mode >> 1) && sg.zb - sg.za + 1 < 2) ok = false;      // two planes to hand over
                ((sg.mode >> 1) ? faces : rest).push_back(sg);
            }
            f2.push_back((int)s2.size());
            s2.insert(s2.end(), faces.begin(), faces.end());
            s2.insert(s2.end(), rest.begin(), rest.end());
        }
        f2.push_back((int)s2.size());
        if (!ok) { SwPlanParams q = p; q.part = 4; return sw_build_plan(q, segs, first); }                   // degenerate cut: separate edge CTAs
        segs.swap(s2); first.swap(f2);
    }
    return p.part;
}
