// solver.cuh -- host-side orchestration of the per-timestep hot path: the launch sequences that
// replace Simulation::linearSolver / diffuse / project / advect / step (simulation.cpp:96-424).
// Every "omp barrier" / loop nest of the reference is a kernel boundary on h->stream.
#pragma once
#include "handle.cuh"
#include "kernels_basic.cuh"
#include "kernels_stencil.cuh"
#include "kernels_reduce.cuh"
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <mutex>
#include <unordered_map>

int fl_halo_exchange(fluid_handle* h, int n, float* const* fields);      // comm.cuh (slab mode)
int fl_lex_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                 float a, float c);                                      // wavefront.cuh
int fl_tiled_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                   float a, float c, bool* handled);                     // tiled.cuh
bool fl_swap_instead_of_copy(fluid_handle* h);                            // sweep_stream.cuh
bool fl_slab_batch4(fluid_handle* h);                                     // sweep_stream.cuh
void fl_tiled_release(fluid_handle* h);                                  // tiled.cuh
void fl_comm_release(fluid_handle* h);                                   // comm.cuh
int fl_halo_exchange_on(fluid_handle* h, int n, float* const* fields, cudaStream_t stream);   // comm.cuh
// comm.cuh: halo hand-over by peer stores + arrival flags (sweeps of the streaming kernel)
bool fl_push_usable(fluid_handle* h);
int fl_halo_push(fluid_handle* h, int n, float* const* fields, cudaStream_t stream, unsigned seq, int nplanes = 2);
int fl_halo_wait(fluid_handle* h, cudaStream_t stream, unsigned seq);
int fl_push_targets(fluid_handle* h, int n, float* const* fields, float* (*peer)[4], unsigned** flag, int* zdst);
bool fl_peer_mapped(fluid_handle* h, int r, const float* local);            // comm.cuh: rank r's copy of this buffer is IPC-mapped
int fl_all_agree(fluid_handle* h, bool mine, bool* all);                   // comm.cuh: collective AND over the ranks
int fl_barrier(fluid_handle* h);                                         // comm.cuh

static inline int fl_rebuild_flags(fluid_handle* h)
{
    if (!h->flags_dirty) return 0;
    k_build_flags<<<fl_grid_1d(h->g.cells), 256, 0, h->stream>>>(h->g, h->field[FLUID_OBS] - FL_ZPAD * h->g.plane,
                                                                 h->cflags - FL_ZPAD * h->g.plane);
    FL_CHECK_LAUNCH(h);
    h->flags_dirty = false;
    ++h->flags_version;
    return 0;
}

// host scalars with the reference's exact float expressions (compiled without contraction)
static inline float fl_diffuse_a(const fluid_handle* h)
{
    // simulation.cpp:282: dt * diff * width * height * depth, left to right, ints -> float
    volatile float a = h->dt * h->diff;
    a = a * (float)h->g.W; a = a * (float)h->g.H; a = a * (float)h->g.D;
    return a;
}
static inline float fl_project_h(const fluid_handle* h)
{
    // simulation.cpp:295: 1.0f / std::cbrt(float(int product))
    return 1.0f / cbrtf((float)(h->g.W * h->g.H * h->g.D));
}

static inline int fl_set_bounds(fluid_handle* h, int b, int fid)
{
    if (int e = fl_rebuild_flags(h)) return e;
    QuadLaunch q = fl_quad_launch(h->g, h->g.Dl);
    k_set_bounds<<<q.grid, q.block, 0, h->stream>>>(h->g, h->field[fid], h->cflags, b);
    FL_CHECK_LAUNCH(h);
    return 0;
}

// linearSolver(b, field, prev, a, c) for up to three fields that share a and c
// (simulation.cpp:251-273): acc sweeps, setBounds after each.
static inline cudaEvent_t fl_prof_event(fluid_handle* h)
{
    cudaEvent_t e = nullptr;
    if (!h->ev_pool.empty()) { e = h->ev_pool.back(); h->ev_pool.pop_back(); }
    else cudaEventCreate(&e);
    return e;
}

static inline int fl_lin_solve_impl(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                                    float a, float c);

static inline int fl_lin_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                               float a, float c)
{
    if (!h->profiling) return fl_lin_solve_impl(h, nb, b, fid, pid, a, c);
    fluid_handle::ProfRec r{ fl_prof_event(h), fl_prof_event(h), (long long)nb * h->acc };
    cudaEventRecord(r.a, h->stream);
    int e = fl_lin_solve_impl(h, nb, b, fid, pid, a, c);
    cudaEventRecord(r.b, h->stream);
    h->prof.push_back(r);
    return e;
}

static inline int fl_lin_solve_impl(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                                    float a, float c)
{
    if (int e = fl_rebuild_flags(h)) return e;
    if (h->acc <= 0) return 0;
    const unsigned solver = h->flags & FLUID_SOLVER_MASK;
    if (solver == FLUID_SOLVER_LEX) return fl_lex_solve(h, nb, b, fid, pid, a, c);
    if (!(h->flags & FLUID_NO_TILED)) {
        bool handled = false;
        if (int e = fl_tiled_solve(h, nb, b, fid, pid, a, c, &handled)) return e;
        if (handled) return 0;
    }
    const float c_recip = 1.0f / c;                                       // simulation.cpp:257
    const Grid& g = h->g;
    QuadLaunch q = fl_quad_launch(g, g.Dl * nb);
    SweepBatch sb;
    float* cur[3]; float* alt[3];
    for (int i = 0; i < nb; ++i) {
        cur[i] = h->field[fid[i]]; alt[i] = h->scratch[i];
        sb.rhs[i] = h->field[pid[i]]; sb.b[i] = b[i];
    }
    if (solver == FLUID_SOLVER_JACOBI) {
        dim3 ge((2 * (g.W + 2) + 2 * g.H + 127) / 128, g.Dl + 2);
        for (int i = 0; i < nb; ++i) {
            k_copy_edges<<<ge, 128, 0, h->stream>>>(g, cur[i], alt[i]);
            FL_CHECK_LAUNCH(h);
        }
    }
    for (int k = 0; k < h->acc; ++k) {                                    // simulation.cpp:258
        if (solver == FLUID_SOLVER_JACOBI) {
            for (int i = 0; i < nb; ++i) { sb.src[i] = cur[i]; sb.dst[i] = alt[i]; }
            k_sweep_jacobi<<<q.grid, q.block, 0, h->stream>>>(g, sb, h->cflags, a, c_recip);
            FL_CHECK_LAUNCH(h);
            for (int i = 0; i < nb; ++i) { float* t = cur[i]; cur[i] = alt[i]; alt[i] = t; }
        } else {  // red-black, in place: colour 0, colour 1, then setBounds
            for (int i = 0; i < nb; ++i) { sb.src[i] = cur[i]; sb.dst[i] = cur[i]; }
            for (int colour = 0; colour < 2; ++colour) {
                // colour 0 also covers a slab's first halo planes (see k_sweep_rb_half)
                const int zf = (colour == 0 && h->rank > 0) ? 0 : 1;
                const int zl = (colour == 0 && h->rank < h->nranks - 1) ? g.Dl + 1 : g.Dl;
                QuadLaunch qh = fl_quad_launch(g, (zl - zf + 1) * nb);
                k_sweep_rb_half<<<qh.grid, qh.block, 0, h->stream>>>(g, sb, a, c_recip, colour, zf, zl - zf + 1);
                FL_CHECK_LAUNCH(h);
            }
            QuadLaunch qb = fl_quad_launch(g, g.Dl);
            for (int i = 0; i < nb; ++i) {
                k_set_bounds<<<qb.grid, qb.block, 0, h->stream>>>(g, cur[i], h->cflags, b[i]);
                FL_CHECK_LAUNCH(h);
            }
        }
        // The in-place red-black kernels read (and, for colour 0, write) the very halo planes the neighbour's next
        // hand-over lands in, with no other hand-over in between: a one-sided peer store could arrive while they still
        // run.  This path keeps the two-sided NCCL exchange; the ping-pong Jacobi form may hand over by peer stores.
        if (h->nranks > 1) {
            if (solver == FLUID_SOLVER_JACOBI) { if (int e = fl_halo_exchange(h, nb, cur)) return e; }
            else                               { if (int e = fl_halo_exchange_on(h, nb, cur, h->stream)) return e; }
        }
    }
    // Jacobi ping-pong: the result may live in the scratch buffer -> swap ownership, no copy
    for (int i = 0; i < nb; ++i) {
        if (cur[i] != h->field[fid[i]]) { h->scratch[i] = h->field[fid[i]]; h->field[fid[i]] = cur[i]; }
    }
    return 0;
}

static inline int fl_diffuse(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid)
{
    const float a = fl_diffuse_a(h);                                      // simulation.cpp:282
    volatile float c = 6.0f * a; c = 1.0f + c;                            // :283, no contraction
    return fl_lin_solve(h, nb, b, fid, pid, a, c);
}

static inline int fl_divergence(fluid_handle* h)
{
    if (int e = fl_rebuild_flags(h)) return e;
    if (h->nranks > 1) { float* f[1] = { h->field[FLUID_VZ] }; if (int e = fl_halo_exchange(h, 1, f)) return e; }
    QuadLaunch q = fl_quad_launch(h->g, h->g.Dl);
    k_divergence<<<q.grid, q.block, 0, h->stream>>>(h->g, h->field[FLUID_VX], h->field[FLUID_VY],
                                                    h->field[FLUID_VZ], h->cflags,
                                                    h->field[FLUID_DIVERGENCE], h->field[FLUID_PRESSURE],
                                                    fl_project_h(h));
    FL_CHECK_LAUNCH(h);
    if (h->nranks > 1) {   // the Poisson sweeps read both on the halo planes
        float* f[2] = { h->field[FLUID_DIVERGENCE], h->field[FLUID_PRESSURE] };
        if (int e = fl_halo_exchange(h, 2, f)) return e;
    }
    return 0;
}

// recip_div.h's exhaustive host check, once per divisor and process
static inline bool fl_recip_div_ok(float s)
{
    static std::mutex mu;
    static std::unordered_map<unsigned, bool> seen;
    unsigned bits; memcpy(&bits, &s, sizeof bits);
    std::lock_guard<std::mutex> lock(mu);
    auto it = seen.find(bits);
    if (it == seen.end()) it = seen.emplace(bits, fl_recip_div_verify(s)).first;
    return it->second;
}

static inline int fl_gradient(fluid_handle* h)
{
    if (int e = fl_rebuild_flags(h)) return e;
    QuadLaunch q = fl_quad_launch(h->g, h->g.Dl);
    // h a power of two (cubic 2^k grids): x / h and x / (2h) are exact scalings, done as multiplications
    const float hh = fl_project_h(h), h2 = 2.0f * hh;                       // :330 (2.0f * h)
    unsigned hb; memcpy(&hb, &hh, sizeof hb);
    const unsigned he = (hb >> 23) & 0xffu;
    static const bool always_div = getenv("FLUID_GRAD_DIV") != nullptr;     // A/B: always divide
    const bool pow2 = (hb & 0x007fffffu) == 0u && he >= 2u && he <= 252u;   // 1/h, 1/(2h) normal and exact
    int mode = FL_GRAD_DIV;
    if (!always_div) mode = pow2 ? FL_GRAD_POW2 : (fl_recip_div_ok(hh) && fl_recip_div_ok(h2)) ? FL_GRAD_RECIP : FL_GRAD_DIV;
    const GradScale sc = { hh, h2, 1.0f / hh, 1.0f / h2 };
    float *vx = h->field[FLUID_VX], *vy = h->field[FLUID_VY], *vz = h->field[FLUID_VZ];
    if (mode == FL_GRAD_POW2)
        k_gradient<FL_GRAD_POW2><<<q.grid, q.block, 0, h->stream>>>(h->g, h->field[FLUID_PRESSURE], h->cflags, vx, vy, vz, sc);
    else if (mode == FL_GRAD_RECIP)
        k_gradient<FL_GRAD_RECIP><<<q.grid, q.block, 0, h->stream>>>(h->g, h->field[FLUID_PRESSURE], h->cflags, vx, vy, vz, sc);
    else
        k_gradient<FL_GRAD_DIV><<<q.grid, q.block, 0, h->stream>>>(h->g, h->field[FLUID_PRESSURE], h->cflags, vx, vy, vz, sc);
    FL_CHECK_LAUNCH(h);
    if (h->nranks > 1) {
        float* f[3] = { h->field[FLUID_VX], h->field[FLUID_VY], h->field[FLUID_VZ] };
        if (int e = fl_halo_exchange(h, 3, f)) return e;
    }
    return 0;
}

static inline int fl_project(fluid_handle* h)                             // simulation.cpp:289-362
{
    if (int e = fl_divergence(h)) return e;
    const int b = 0, f = FLUID_PRESSURE, p = FLUID_DIVERGENCE;
    if (int e = fl_lin_solve(h, 1, &b, &f, &p, 1.0f, 6.0f)) return e;     // :320
    return fl_gradient(h);
}

int fl_gather_sources(fluid_handle* h, int prev_id, GatherSrc* out);      // fluid_cuda.cu

static inline int fl_advect(fluid_handle* h, int b, int fid, int pid)     // simulation.cpp:367-424
{
    if (int e = fl_rebuild_flags(h)) return e;
    GatherSrc src;
    if (int e = fl_gather_sources(h, pid, &src)) return e;
    if (int e = fl_barrier(h)) return e;          // every rank's slab of the gather source is complete
    static const bool quad = getenv("FLUID_ADVECT_QUAD") != nullptr;        // A/B: the one-thread-per-quad form
    if (quad || h->g.cells >= (1ll << 31)) {
        QuadLaunch q = fl_quad_launch(h->g, h->g.Dl);
        k_advect<<<q.grid, q.block, 0, h->stream>>>(h->g, b, h->field[fid], h->field[pid], src,
                                                    h->field[FLUID_VX], h->field[FLUID_VY],
                                                    h->field[FLUID_VZ], h->cflags, h->dt);
    } else {
        static const int ny_env = getenv("FLUID_ADVECT_NY") ? atoi(getenv("FLUID_ADVECT_NY")) : 0;
        const int NY = (ny_env == 1 || ny_env == 2 || ny_env == 4) ? ny_env : 2;            // rows per thread (2: measured best, 86 vs 92 us at 256^3)
        int bx = 32;
        while (bx < h->g.W && bx < 128) bx <<= 1;
        const dim3 block(bx, 256 / bx, 1);
        const dim3 grid((h->g.W + bx - 1) / bx, (h->g.H + block.y * NY - 1) / (block.y * NY), h->g.Dl);
#define FL_ADV_LAUNCH(SLAB, N) k_advect_rows<SLAB, N><<<grid, block, 0, h->stream>>>(h->g, b, h->field[fid], h->field[pid], src, \
                                   h->field[FLUID_VX], h->field[FLUID_VY], h->field[FLUID_VZ], h->cflags, h->dt)
        if (src.n > 1) { if (NY == 1) FL_ADV_LAUNCH(true, 1); else if (NY == 2) FL_ADV_LAUNCH(true, 2); else FL_ADV_LAUNCH(true, 4); }
        else           { if (NY == 1) FL_ADV_LAUNCH(false, 1); else if (NY == 2) FL_ADV_LAUNCH(false, 2); else FL_ADV_LAUNCH(false, 4); }
#undef FL_ADV_LAUNCH
    }
    FL_CHECK_LAUNCH(h);
    if (h->nranks > 1) {
        float* f[1] = { h->field[fid] };
        if (int e = fl_halo_exchange(h, 1, f)) return e;
    }
    return 0;
}

// advect(1, v_x, v_x_prev); advect(2, v_y, v_y_prev); advect(3, v_z, v_z_prev) (simulation.cpp:125-127) as one
// launch (k_advect3); FLUID_ADVECT_FUSED=0 keeps the three launches (A/B)
static inline int fl_advect_velocity(fluid_handle* h)
{
    static const bool fused_off = getenv("FLUID_ADVECT_FUSED") && atoi(getenv("FLUID_ADVECT_FUSED")) == 0;
    static const bool quad = getenv("FLUID_ADVECT_QUAD") != nullptr;
    if (fused_off || quad || h->g.cells >= (1ll << 31)) {
        if (int e = fl_advect(h, 1, FLUID_VX, FLUID_VX_PREV)) return e;
        if (int e = fl_advect(h, 2, FLUID_VY, FLUID_VY_PREV)) return e;
        return fl_advect(h, 3, FLUID_VZ, FLUID_VZ_PREV);
    }
    if (int e = fl_rebuild_flags(h)) return e;
    const int pid[3] = { FLUID_VX_PREV, FLUID_VY_PREV, FLUID_VZ_PREV };
    GatherSrc3 src;
    for (int c = 0; c < 3; ++c) {
        GatherSrc one = {};
        if (int e = fl_gather_sources(h, pid[c], &one)) return e;
        for (int r = 0; r < 8; ++r) { src.base[c][r] = one.base[r]; src.zlo[r] = one.zlo[r]; src.dl[r] = one.dl[r]; }
        src.n = one.n;
    }
    if (int e = fl_barrier(h)) return e;          // every rank's slabs of the gather sources are complete
    static const int ny_env = getenv("FLUID_ADVECT_NY") ? atoi(getenv("FLUID_ADVECT_NY")) : 0;
    const int NY = (ny_env == 1 || ny_env == 2) ? ny_env : 2;
    int bx = 32;
    while (bx < h->g.W && bx < 128) bx <<= 1;
    const dim3 block(bx, 256 / bx, 1);
    const dim3 grid((h->g.W + bx - 1) / bx, (h->g.H + block.y * NY - 1) / (block.y * NY), h->g.Dl);
#define FL_ADV3_LAUNCH(SLAB, N) k_advect3<SLAB, N><<<grid, block, 0, h->stream>>>(h->g, h->field[FLUID_VX], h->field[FLUID_VY], \
                                    h->field[FLUID_VZ], h->field[FLUID_VX_PREV], h->field[FLUID_VY_PREV], h->field[FLUID_VZ_PREV], \
                                    src, h->cflags, h->dt)
    if (src.n > 1) { if (NY == 1) FL_ADV3_LAUNCH(true, 1); else FL_ADV3_LAUNCH(true, 2); }
    else           { if (NY == 1) FL_ADV3_LAUNCH(false, 1); else FL_ADV3_LAUNCH(false, 2); }
#undef FL_ADV3_LAUNCH
    FL_CHECK_LAUNCH(h);
    if (h->nranks > 1) {
        float* f[3] = { h->field[FLUID_VX], h->field[FLUID_VY], h->field[FLUID_VZ] };
        if (int e = fl_halo_exchange(h, 3, f)) return e;
    }
    return 0;
}

static inline int fl_inflow_density(fluid_handle* h)                      // simulation.cpp:65-67
{
    dim3 grid((h->g.H + 127) / 128, h->g.Dl + 2 * FL_HALO);
    k_inflow_density<<<grid, 128, 0, h->stream>>>(h->g, h->field[FLUID_DENS], 0.001f);
    FL_CHECK_LAUNCH(h);
    return 0;
}

static inline int fl_inflow_velocity(fluid_handle* h)                     // simulation.cpp:103-105
{
    dim3 grid((h->g.H + 127) / 128, h->g.Dl + 2 * FL_HALO);
    k_inflow_velocity<<<grid, 128, 0, h->stream>>>(h->g, h->field[FLUID_VX], h->field[FLUID_VY],
                                                   h->field[FLUID_VZ], (float)h->speed);
    FL_CHECK_LAUNCH(h);
    return 0;
}

static inline int fl_copy_field(fluid_handle* h, int dst, int src)
{
    FL_CUDA(h, cudaMemcpyAsync(h->field[dst] - FL_ZPAD * h->g.plane, h->field[src] - FL_ZPAD * h->g.plane, (size_t)h->g.cells * sizeof(float),
                               cudaMemcpyDeviceToDevice, h->stream));
    return 0;
}

// Simulation::step() without the frame write (simulation.cpp:96-138)
static inline int fl_step_only(fluid_handle* h)
{
    if (int e = fl_inflow_velocity(h)) return e;                          // :103-105
    // The density diffusion of :135 reads only dens and buffer, which nothing between here and :135 touches, uses
    // the same a and c as the velocity diffusions (:282-283) and its result is not read before :136.  On the
    // streaming single-GPU path it therefore rides along as a fourth field of the batched velocity solve: same
    // arithmetic, same results, 80 launches fewer and longer z-ranges per CTA (4 extra planes per range of ~130
    // instead of per ~33).
    static const bool no_batch4 = getenv("FLUID_NO_BATCH4") != nullptr;
    const bool swap = fl_swap_instead_of_copy(h);
    // z-slabs batch the fourth field too (its ping-pong partner is IPC-mapped like the others); they keep the copies
    const bool batch4 = (swap || fl_slab_batch4(h)) && !no_batch4 && !(h->flags & FLUID_ELIDE_DEAD);
    if (swap) {
        // :108-110 v_prev = v without moving bytes: the buffers trade places (v_prev now holds v, v a stale buffer
        // that the out-of-place solve below overwrites completely) and the solve's first sweep reads its initial
        // guess -- which the copy would have made equal to v_prev -- from v_prev.  3 x 70 MB less traffic at 256^3.
        for (int k = 0; k < 3; ++k) { float* t = h->field[FLUID_VX + k]; h->field[FLUID_VX + k] = h->field[FLUID_VX_PREV + k]; h->field[FLUID_VX_PREV + k] = t; }
        h->x_from_rhs = 0x7u;
    } else {
        if (int e = fl_copy_field(h, FLUID_VX_PREV, FLUID_VX)) return e;  // :108-110
        if (int e = fl_copy_field(h, FLUID_VY_PREV, FLUID_VY)) return e;
        if (int e = fl_copy_field(h, FLUID_VZ_PREV, FLUID_VZ)) return e;
    }
    {   // :115-117 the three velocity diffusions share a and c -> one batched launch per sweep
        const int b[4] = { 1, 2, 3, 0 }, f[4] = { FLUID_VX, FLUID_VY, FLUID_VZ, FLUID_DENS },
                  p[4] = { FLUID_VX_PREV, FLUID_VY_PREV, FLUID_VZ_PREV, FLUID_BUFFER };
        if (int e = fl_diffuse(h, batch4 ? 4 : 3, b, f, p)) return e;
    }
    if (int e = fl_project(h)) return e;                                  // :120
    if (int e = fl_advect_velocity(h)) return e;                          // :125-127, in this order
    if (int e = fl_project(h)) return e;                                  // :130
    if (!(h->flags & FLUID_ELIDE_DEAD) && !batch4) {                      // :135 (dead work, F2)
        const int b = 0, f = FLUID_DENS, p = FLUID_BUFFER;
        if (int e = fl_diffuse(h, 1, &b, &f, &p)) return e;
    }
    if (int e = fl_advect(h, 0, FLUID_DENS, FLUID_BUFFER)) return e;      // :136
    return fl_barrier(h);        // slab mode: nobody overwrites a gather source while a peer still reads it
}

// one iteration of run()'s loop body (simulation.cpp:65-71)
static inline int fl_run_iteration(fluid_handle* h)
{
    if (int e = fl_inflow_density(h)) return e;                           // :65-67
    if (int e = fl_copy_field(h, FLUID_BUFFER, FLUID_DENS)) return e;     // :70
    return fl_step_only(h);
}
