/* oracle/fluid_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference hot path (Simulation::step() and its
 * callees, /root/reference/simulation.cpp:96-424) used ONLY as the parity
 * checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.
 * The product (libfluid_cuda) never links, loads or calls this.
 *
 * Parity status: PINNED.  With order = FO_ORDER_LEX every function here is
 * checked bit-for-bit against the unmodified reference object code
 * (oracle/_ref/libfluid_ref.so, built by oracle/Makefile) in
 * tests/test_oracle_vs_reference.py, and against committed golden vectors that
 * were produced by that reference build (tests/golden/make_golden.py).
 *
 * FO_ORDER_RB / FO_ORDER_JACOBI are the "ordering-matched CPU variants" of
 * linearSolver: identical update formula, identical "sweep solids then zero
 * them in setBounds" rule, identical setBounds after every sweep -- only the
 * visiting order differs (the reference's in-place lexicographic order cannot
 * be run in parallel; its own OpenMP build races).
 */
#ifndef FLUID_ORACLE_H
#define FLUID_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

enum { FO_VX = 0, FO_VY, FO_VZ, FO_VX_PREV, FO_VY_PREV, FO_VZ_PREV,
       FO_DENS, FO_BUFFER, FO_OBS, FO_PRESSURE, FO_DIVERGENCE, FO_NFIELDS };

enum { FO_ORDER_LEX = 0,     /* reference order: x outer, y, z inner, in place (simulation.cpp:260-262) */
       FO_ORDER_RB = 1,      /* red-black in place: colour (i+j+l)&1 == 0 first, then == 1            */
       FO_ORDER_JACOBI = 2   /* all cells from the previous sweep's (post-setBounds) values            */ };

typedef struct fo_sim {
    int width, height, depth;        /* simulation.h:44 */
    int speed;                       /* simulation.h:47 */
    float dt, diff, visc;            /* simulation.h:50-52 */
    int acc;                         /* simulation.h:53 */
    int order;                       /* FO_ORDER_* */
    int elide_dead;                  /* 1: skip the dead density diffusion */
    int size;                        /* simulation.cpp:35 */
    float* f[FO_NFIELDS];            /* simulation.h:16-27, reference layout x + y*Wp + z*Wp*Hp */
    float* scratch;                  /* Jacobi source copy */
} fo_sim;

fo_sim* fo_create(int w, int h, int d, int speed, float dt, float diff, float visc, int acc, int order);
void    fo_destroy(fo_sim* s);
float*  fo_field(fo_sim* s, int id);
int     fo_size(const fo_sim* s);
void    fo_set_order(fo_sim* s, int order);
void    fo_set_acc(fo_sim* s, int acc);
void    fo_set_elide_dead(fo_sim* s, int on);

void fo_add_obstacle(fo_sim* s, int x, int y, int z);                       /* simulation.cpp:155 */
void fo_add_density(fo_sim* s, int x, int y, int z, float amount);           /* simulation.cpp:163 */
void fo_set_velocity(fo_sim* s, int x, int y, int z, float ax, float ay, float az); /* :171 */

void fo_set_bounds(fo_sim* s, int b, int fid);                               /* simulation.cpp:183 */
void fo_linear_solver(fo_sim* s, int b, int fid, int prev_id, float a, float c); /* :251 */
void fo_diffuse(fo_sim* s, int b, int fid, int prev_id);                     /* simulation.cpp:278 */
void fo_divergence(fo_sim* s);                                               /* simulation.cpp:295-319 */
void fo_gradient(fo_sim* s);                                                 /* simulation.cpp:322-361 */
void fo_project(fo_sim* s);                                                  /* simulation.cpp:289 */
void fo_advect(fo_sim* s, int b, int fid, int prev_id);                      /* simulation.cpp:367 */
void fo_inflow_density(fo_sim* s);                                           /* simulation.cpp:65-67 */
void fo_inflow_velocity(fo_sim* s);                                          /* simulation.cpp:103-105 */
void fo_step(fo_sim* s);                                                     /* simulation.cpp:96-150 */
void fo_run_iteration(fo_sim* s);                                            /* simulation.cpp:65-71 */
void fo_run_iterations(fo_sim* s, int n);

/* host scalars computed with the reference's exact float expressions */
float fo_diffuse_a(const fo_sim* s);                                         /* simulation.cpp:282 */
float fo_project_h(const fo_sim* s);                                         /* simulation.cpp:295 */

#ifdef __cplusplus
}
#endif
#endif
