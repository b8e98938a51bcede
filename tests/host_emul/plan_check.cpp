// plan_check.cpp -- TEST INFRASTRUCTURE.  The streaming sweep's work decomposition (fluid_simulation_b200/csrc/sweep_plan.h,
// pure host code) behind a C entry point, so that tests/test_sweep_plan.py can check its properties for any shape.
// Built by the test with: g++ -std=c++17 -O1 -shared -fPIC
#include "../../fluid_simulation_b200/csrc/sweep_plan.h"
#include <cstring>

// params: Dl, H, nb, ty, cps, part, sm_count, lo_face, hi_face, edge, late_extra, zc_override, plan_mode, plan_min_fill,
//         last_w32, segcost
// Returns the part that was built; *n_segs / *n_ctas receive the sizes; segs (4 ints each) and first (n_ctas + 1 ints)
// are filled up to the given capacities.
extern "C" int plan_build(const int* prm, int cap_segs, int cap_first, int* n_segs, int* n_ctas, int* segs_out, int* first_out)
{
    SwPlanParams p;
    p.Dl = prm[0]; p.H = prm[1]; p.nb = prm[2]; p.ty = prm[3]; p.cps = prm[4]; p.part = prm[5]; p.segcost = SW_SEGCOST;
    p.sm_count = prm[6]; p.lo_face = prm[7] != 0; p.hi_face = prm[8] != 0; p.edge = prm[9]; p.late_extra = prm[10];
    p.zc_override = prm[11]; p.plan_mode = prm[12]; p.plan_min_fill = prm[13]; p.last_w32 = prm[14]; p.segcost = prm[15];
    std::vector<SwSeg> segs; std::vector<int> first;
    const int built = sw_build_plan(p, segs, first);
    *n_segs = (int)segs.size(); *n_ctas = (int)first.size() - 1;
    if ((int)segs.size() <= cap_segs) memcpy(segs_out, segs.data(), segs.size() * sizeof(SwSeg));
    if ((int)first.size() <= cap_first) memcpy(first_out, first.data(), first.size() * sizeof(int));
    return built;
}
