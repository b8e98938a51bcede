// oracle/ref_harness.cpp  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin extern "C" shim around the UNMODIFIED reference `Simulation` class.  The
// reference sources are compiled where they lie (/root/reference/simulation.cpp,
// object_loader.cpp; recipe in oracle/Makefile) and linked into
// oracle/_ref/libfluid_ref.so together with this file.  Nothing from the
// reference is copied into the repo.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// `--impl reference` legs may load the resulting library, and only as the
// checker / the timed CPU baseline -- never on the product path.
//
// The reference keeps its fields and per-phase methods private
// (simulation.h:12-40).  We reach them with the classic private->public include
// trick; the class layout is unaffected because access
// specifiers do not reorder members within one class for GCC.
#include <vector>
#include <fstream>
#include <cstring>
#include <omp.h>

#define private public
#include "simulation.h"          // from -I/root/reference
#undef private

extern "C" {

// Field ids shared with include/fluid_cuda.h and oracle/fluid_oracle.h.
static std::vector<float>* field_of(Simulation* s, int id) {
    switch (id) {
        case 0:  return &s->v_x;        case 1:  return &s->v_y;      case 2: return &s->v_z;
        case 3:  return &s->v_x_prev;   case 4:  return &s->v_y_prev; case 5: return &s->v_z_prev;
        case 6:  return &s->dens;       case 7:  return &s->buffer;   case 8: return &s->obs;
        case 9:  return &s->pressure;   case 10: return &s->divergence;
    }
    return nullptr;
}

void* ref_create(int w, int h, int d, int iter, int speed, float dt, float diff, float visc, int acc) {
    return new Simulation(w, h, d, iter, speed, dt, diff, visc, acc);   // simulation.cpp:17
}
void ref_destroy(void* p) { delete static_cast<Simulation*>(p); }

int ref_size(void* p) { return static_cast<Simulation*>(p)->size; }     // simulation.cpp:35
float* ref_field(void* p, int id) {
    auto* f = field_of(static_cast<Simulation*>(p), id);
    return f ? f->data() : nullptr;
}
void ref_set_acc(void* p, int acc) { static_cast<Simulation*>(p)->acc = acc; }
void ref_set_threads(int n) { omp_set_num_threads(n); }
int  ref_max_threads(void) { return omp_get_max_threads(); }

// public mutators, simulation.cpp:155-178
void ref_add_obstacle(void* p, int x, int y, int z) { static_cast<Simulation*>(p)->addObstacle(x, y, z); }
void ref_add_density(void* p, int x, int y, int z, float a) { static_cast<Simulation*>(p)->addDensity(x, y, z, a); }
void ref_set_velocity(void* p, int x, int y, int z, float ax, float ay, float az) {
    static_cast<Simulation*>(p)->setVelocity(x, y, z, ax, ay, az);
}

// private phases, called outside any parallel region -> orphaned `omp for`s run
// on one thread, deterministic.
void ref_set_bounds(void* p, int b, int fid) {
    auto* s = static_cast<Simulation*>(p); s->setBounds(b, *field_of(s, fid));          // simulation.cpp:183
}
void ref_linear_solver(void* p, int b, int fid, int prev_id, float a, float c) {
    auto* s = static_cast<Simulation*>(p);
    s->linearSolver(b, *field_of(s, fid), *field_of(s, prev_id), a, c);                 // simulation.cpp:251
}
void ref_diffuse(void* p, int b, int fid, int prev_id) {
    auto* s = static_cast<Simulation*>(p); s->diffuse(b, *field_of(s, fid), *field_of(s, prev_id));   // :278
}
void ref_project(void* p) {
    auto* s = static_cast<Simulation*>(p);
    s->project(s->v_x, s->v_y, s->v_z, s->pressure, s->divergence);                     // simulation.cpp:289
}
void ref_advect(void* p, int b, int fid, int prev_id) {
    auto* s = static_cast<Simulation*>(p); s->advect(b, *field_of(s, fid), *field_of(s, prev_id));    // :367
}

// step() opens its own omp team (simulation.cpp:98): thread count = omp_set_num_threads.
// The ofstreams are never opened, so the frame writes at :143-147 are no-ops.
void ref_step(void* p) { static_cast<Simulation*>(p)->step(); }

// One iteration of run()'s loop body, simulation.cpp:65-71 (inlet density on the
// x=1 face through the public mutator, buffer = dens, step()).
void ref_run_iteration(void* p) {
    auto* s = static_cast<Simulation*>(p);
    for (int j = 1; j <= s->height; ++j)
        for (int k = 1; k <= s->depth; ++k)
            s->addDensity(1, j, k, 0.001f);
    s->buffer = s->dens;
    s->step();
}
void ref_run_iterations(void* p, int n) { for (int i = 0; i < n; ++i) ref_run_iteration(p); }

// The reference STL voxeliser (object_loader.cpp:270); mask is NOT reproducible
// run to run (thread-id seeded RNG), so callers dump the mask and share the bytes.
void ref_load_stl(void* p, const char* path, float scale, float rx, float ry, float rz,
                  float tx, float ty, float tz) {
    loadSTLIntoObstacles(path, *static_cast<Simulation*>(p), scale, rx, ry, rz, tx, ty, tz);
}

}  // extern "C"
