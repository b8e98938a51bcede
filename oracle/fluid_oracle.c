/* oracle/fluid_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see fluid_oracle.h).
 *
 * CPU restatement of /root/reference/simulation.cpp:96-424.  Every floating
 * point expression keeps the reference's operand order and is compiled with
 * -ffp-contract=off so that no FMA is formed (the reference build has none,
 * SURVEY.md H3).  Loops whose visiting order cannot change the result are run
 * z-outer / x-inner for cache friendliness; the Gauss-Seidel sweep keeps the
 * reference's x-outer / z-inner order when order == FO_ORDER_LEX.
 */
#include "fluid_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define WP(s) ((s)->width + 2)
#define HP(s) ((s)->height + 2)
/* simulation.h:9 global_at */
#define AT(s, x, y, z) ((size_t)(x) + (size_t)(y) * WP(s) + (size_t)(z) * WP(s) * HP(s))

fo_sim* fo_create(int w, int h, int d, int speed, float dt, float diff, float visc, int acc, int order)
{
    fo_sim* s = (fo_sim*)calloc(1, sizeof(fo_sim));
    if (!s) return NULL;
    s->width = w; s->height = h; s->depth = d; s->speed = speed;
    s->dt = dt; s->diff = diff; s->visc = visc; s->acc = acc; s->order = order;
    s->elide_dead = 0;
    s->size = (w + 2) * (h + 2) * (d + 2);                     /* simulation.cpp:35 */
    for (int i = 0; i < FO_NFIELDS; ++i)                       /* simulation.cpp:38-43: zero filled */
        s->f[i] = (float*)calloc((size_t)s->size, sizeof(float));
    s->scratch = (float*)calloc((size_t)s->size, sizeof(float));
    return s;
}

void fo_destroy(fo_sim* s)
{
    if (!s) return;
    for (int i = 0; i < FO_NFIELDS; ++i) free(s->f[i]);
    free(s->scratch);
    free(s);
}

float* fo_field(fo_sim* s, int id) { return (id >= 0 && id < FO_NFIELDS) ? s->f[id] : NULL; }
int    fo_size(const fo_sim* s) { return s->size; }
void   fo_set_order(fo_sim* s, int order) { s->order = order; }
void   fo_set_acc(fo_sim* s, int acc) { s->acc = acc; }
void   fo_set_elide_dead(fo_sim* s, int on) { s->elide_dead = on; }

void fo_add_obstacle(fo_sim* s, int x, int y, int z) { s->f[FO_OBS][AT(s, x, y, z)] = 1.0f; }
void fo_add_density(fo_sim* s, int x, int y, int z, float amount) { s->f[FO_DENS][AT(s, x, y, z)] += amount; }
void fo_set_velocity(fo_sim* s, int x, int y, int z, float ax, float ay, float az)
{
    size_t p = AT(s, x, y, z);
    s->f[FO_VX][p] = ax; s->f[FO_VY][p] = ay; s->f[FO_VZ][p] = az;
}

/* ---- setBounds, simulation.cpp:183-246 ------------------------------------
 * order of phases matters: faces copy the PRE-zero interior value, then solids
 * are zeroed, then (velocity only) fluid cells touching a solid are zeroed.
 * Edge and corner ghosts are never written. */
void fo_set_bounds(fo_sim* s, int b, int fid)
{
    float* f = s->f[fid];
    const float* obs = s->f[FO_OBS];
    const int W = s->width, H = s->height, D = s->depth;

    for (int z = 1; z <= D; ++z)                                           /* :187-192 x faces */
        for (int y = 1; y <= H; ++y) {
            f[AT(s, 0, y, z)]     = (b == 1) ? -f[AT(s, 1, y, z)] : f[AT(s, 1, y, z)];
            f[AT(s, W + 1, y, z)] = f[AT(s, W, y, z)];                     /* outflow copy */
        }
    for (int z = 1; z <= D; ++z)                                           /* :196-202 y faces */
        for (int x = 1; x <= W; ++x) {
            f[AT(s, x, 0, z)]     = (b == 2) ? -f[AT(s, x, 1, z)] : f[AT(s, x, 1, z)];
            f[AT(s, x, H + 1, z)] = (b == 2) ? -f[AT(s, x, H, z)] : f[AT(s, x, H, z)];
        }
    for (int y = 1; y <= H; ++y)                                           /* :206-215 z faces */
        for (int x = 1; x <= W; ++x) {
            f[AT(s, x, y, 0)]     = (b == 3) ? -f[AT(s, x, y, 1)] : f[AT(s, x, y, 1)];
            f[AT(s, x, y, D + 1)] = (b == 3) ? -f[AT(s, x, y, D)] : f[AT(s, x, y, D)];
        }
    for (int z = 1; z <= D; ++z)                                           /* :219-223 solids -> 0 */
        for (int y = 1; y <= H; ++y)
            for (int x = 1; x <= W; ++x)
                if (obs[AT(s, x, y, z)] == 1.0f) f[AT(s, x, y, z)] = 0.0f;

    if (b == 1 || b == 2 || b == 3)                                        /* :227-245 no-slip */
        for (int z = 1; z <= D; ++z)
            for (int y = 1; y <= H; ++y)
                for (int x = 1; x <= W; ++x) {
                    if (obs[AT(s, x, y, z)] == 1.0f) continue;
                    int adj = (x + 1 <= W && obs[AT(s, x + 1, y, z)] == 1.0f) ||
                              (x - 1 >= 1 && obs[AT(s, x - 1, y, z)] == 1.0f) ||
                              (y + 1 <= H && obs[AT(s, x, y + 1, z)] == 1.0f) ||
                              (y - 1 >= 1 && obs[AT(s, x, y - 1, z)] == 1.0f) ||
                              (z + 1 <= D && obs[AT(s, x, y, z + 1)] == 1.0f) ||
                              (z - 1 >= 1 && obs[AT(s, x, y, z - 1)] == 1.0f);
                    if (adj) f[AT(s, x, y, z)] = 0.0f;
                }
}

/* the 7-point update of simulation.cpp:263-269, neighbour sum strictly left to right */
#define RELAX(dst, src, p0, i, j, l)                                                         \
    (dst)[AT(s, i, j, l)] =                                                                  \
        ((p0)[AT(s, i, j, l)] +                                                              \
         a * ((src)[AT(s, (i) + 1, j, l)] + (src)[AT(s, (i) - 1, j, l)] +                    \
              (src)[AT(s, i, (j) + 1, l)] + (src)[AT(s, i, (j) - 1, l)] +                    \
              (src)[AT(s, i, j, (l) + 1)] + (src)[AT(s, i, j, (l) - 1)])) * c_recip

/* ---- linearSolver, simulation.cpp:251-273 ---------------------------------
 * acc sweeps over ALL interior cells (solids included, no obs test), setBounds
 * after every sweep. */
void fo_linear_solver(fo_sim* s, int b, int fid, int prev_id, float a, float c)
{
    float* f = s->f[fid];
    const float* p0 = s->f[prev_id];
    const int W = s->width, H = s->height, D = s->depth;
    const float c_recip = 1.0f / c;                                        /* :257 */

    for (int k = 0; k < s->acc; ++k) {                                     /* :258 */
        if (s->order == FO_ORDER_LEX) {
            for (int i = 1; i <= W; ++i)                                   /* :260-262 visiting order */
                for (int j = 1; j <= H; ++j)
                    for (int l = 1; l <= D; ++l)
                        RELAX(f, f, p0, i, j, l);
        } else if (s->order == FO_ORDER_RB) {
            for (int colour = 0; colour < 2; ++colour)
                for (int l = 1; l <= D; ++l)
                    for (int j = 1; j <= H; ++j)
                        for (int i = 1 + ((1 + j + l + colour) & 1); i <= W; i += 2)
                            RELAX(f, f, p0, i, j, l);
        } else { /* FO_ORDER_JACOBI */
            float* old = s->scratch;
            memcpy(old, f, (size_t)s->size * sizeof(float));
            for (int l = 1; l <= D; ++l)
                for (int j = 1; j <= H; ++j)
                    for (int i = 1; i <= W; ++i)
                        RELAX(f, old, p0, i, j, l);
        }
        fo_set_bounds(s, b, fid);                                          /* :271 */
    }
}

float fo_diffuse_a(const fo_sim* s)
{
    /* simulation.cpp:282: float * float * int * int * int, left to right, ints converted to float */
    return s->dt * s->diff * s->width * s->height * s->depth;
}

float fo_project_h(const fo_sim* s)
{
    /* simulation.cpp:295: cbrt of the INT product converted to float (float overload -> cbrtf) */
    return 1.0f / cbrtf((float)(s->width * s->height * s->depth));
}

void fo_diffuse(fo_sim* s, int b, int fid, int prev_id)                    /* simulation.cpp:278-284 */
{
    float a = fo_diffuse_a(s);
    fo_linear_solver(s, b, fid, prev_id, a, 1.0f + 6.0f * a);
}

/* ---- project part 1: divergence + p = 0, simulation.cpp:295-319 ---------- */
void fo_divergence(fo_sim* s)
{
    const float* vx = s->f[FO_VX]; const float* vy = s->f[FO_VY]; const float* vz = s->f[FO_VZ];
    const float* obs = s->f[FO_OBS];
    float* p = s->f[FO_PRESSURE]; float* div = s->f[FO_DIVERGENCE];
    const int W = s->width, H = s->height, D = s->depth;
    const float h = fo_project_h(s);

    for (int k = 1; k <= D; ++k)
        for (int j = 1; j <= H; ++j)
            for (int i = 1; i <= W; ++i) {
                size_t c = AT(s, i, j, k);
                if (obs[c] == 1.0f) { div[c] = 0.0f; p[c] = 0.0f; continue; }          /* :301-305 */
                float acc = 0.0f;                                                      /* :306-312 */
                if (i + 1 <= W && obs[AT(s, i + 1, j, k)] == 0.0f) acc += vx[AT(s, i + 1, j, k)];
                if (i - 1 >= 1 && obs[AT(s, i - 1, j, k)] == 0.0f) acc -= vx[AT(s, i - 1, j, k)];
                if (j + 1 <= H && obs[AT(s, i, j + 1, k)] == 0.0f) acc += vy[AT(s, i, j + 1, k)];
                if (j - 1 >= 1 && obs[AT(s, i, j - 1, k)] == 0.0f) acc -= vy[AT(s, i, j - 1, k)];
                if (k + 1 <= D && obs[AT(s, i, j, k + 1)] == 0.0f) acc += vz[AT(s, i, j, k + 1)];
                if (k - 1 >= 1 && obs[AT(s, i, j, k - 1)] == 0.0f) acc -= vz[AT(s, i, j, k - 1)];
                div[c] = -0.5f * h * acc;                                              /* :314 */
                p[c] = 0.0f;
            }
    fo_set_bounds(s, 0, FO_DIVERGENCE);                                                /* :318 */
    fo_set_bounds(s, 0, FO_PRESSURE);                                                  /* :319 */
}

/* one-sided / central pressure difference along one axis, simulation.cpp:328-336 etc. */
static float axis_gradient(int plus_open, int minus_open, float pp, float pc, float pm, float h)
{
    if (plus_open && minus_open) return (pp - pm) / (2.0f * h);
    if (plus_open)               return (pp - pc) / h;
    if (minus_open)              return (pc - pm) / h;
    return 0.0f;
}

/* ---- project part 3: gradient subtraction, simulation.cpp:322-361 -------- */
void fo_gradient(fo_sim* s)
{
    float* vx = s->f[FO_VX]; float* vy = s->f[FO_VY]; float* vz = s->f[FO_VZ];
    const float* obs = s->f[FO_OBS]; const float* p = s->f[FO_PRESSURE];
    const int W = s->width, H = s->height, D = s->depth;
    const float h = fo_project_h(s);

    for (int k = 1; k <= D; ++k)
        for (int j = 1; j <= H; ++j)
            for (int i = 1; i <= W; ++i) {
                size_t c = AT(s, i, j, k);
                if (obs[c] == 1.0f) continue;                                          /* :326 */
                int xp = i + 1 <= W && obs[AT(s, i + 1, j, k)] == 0.0f;
                int xm = i - 1 >= 1 && obs[AT(s, i - 1, j, k)] == 0.0f;
                int yp = j + 1 <= H && obs[AT(s, i, j + 1, k)] == 0.0f;
                int ym = j - 1 >= 1 && obs[AT(s, i, j - 1, k)] == 0.0f;
                int zp = k + 1 <= D && obs[AT(s, i, j, k + 1)] == 0.0f;
                int zm = k - 1 >= 1 && obs[AT(s, i, j, k - 1)] == 0.0f;
                vx[c] -= axis_gradient(xp, xm, p[AT(s, i + 1, j, k)], p[c], p[AT(s, i - 1, j, k)], h);
                vy[c] -= axis_gradient(yp, ym, p[AT(s, i, j + 1, k)], p[c], p[AT(s, i, j - 1, k)], h);
                vz[c] -= axis_gradient(zp, zm, p[AT(s, i, j, k + 1)], p[c], p[AT(s, i, j, k - 1)], h);
            }
    fo_set_bounds(s, 1, FO_VX);                                                        /* :359-361 */
    fo_set_bounds(s, 2, FO_VY);
    fo_set_bounds(s, 3, FO_VZ);
}

void fo_project(fo_sim* s)                                                             /* :289-362 */
{
    fo_divergence(s);
    fo_linear_solver(s, 0, FO_PRESSURE, FO_DIVERGENCE, 1.0f, 6.0f);                    /* :320 */
    fo_gradient(s);
}

static float clampf(float v, float lo, float hi) { return (v < lo) ? lo : (hi < v) ? hi : v; } /* std::clamp */

/* ---- advect, simulation.cpp:367-424 -------------------------------------- */
void fo_advect(fo_sim* s, int b, int fid, int prev_id)
{
    float* f = s->f[fid];
    const float* prev = s->f[prev_id];
    const float* vxf = s->f[FO_VX]; const float* vyf = s->f[FO_VY]; const float* vzf = s->f[FO_VZ];
    const float* obs = s->f[FO_OBS];
    const int W = s->width, H = s->height, D = s->depth;
    const float dt = s->dt;

    for (int k = 1; k <= D; ++k)
        for (int j = 1; j <= H; ++j)
            for (int i = 1; i <= W; ++i) {
                size_t c = AT(s, i, j, k);
                if (obs[c] == 1.0f) { f[c] = 0.0f; continue; }                         /* :375-378 */
                float vx = (b == 1) ? prev[c] : vxf[c];                                /* :380-382 */
                float vy = (b == 2) ? prev[c] : vyf[c];
                float vz = (b == 3) ? prev[c] : vzf[c];
                float xb = i - dt * W * vx;                                            /* :384-386 */
                float yb = j - dt * H * vy;
                float zb = k - dt * D * vz;
                xb = clampf(xb, 0.5f, (float)W + 0.5f);                                /* :388-390 */
                yb = clampf(yb, 0.5f, (float)H + 0.5f);
                zb = clampf(zb, 0.5f, (float)D + 0.5f);
                int i0 = (int)floorf(xb), i1 = i0 + 1;                                 /* :392-397 */
                int j0 = (int)floorf(yb), j1 = j0 + 1;
                int k0 = (int)floorf(zb), k1 = k0 + 1;
                float sx = xb - i0, sy = yb - j0, sz = zb - k0;                        /* :399-401 */
                float c000 = prev[AT(s, i0, j0, k0)], c100 = prev[AT(s, i1, j0, k0)];  /* :403-410 */
                float c010 = prev[AT(s, i0, j1, k0)], c110 = prev[AT(s, i1, j1, k0)];
                float c001 = prev[AT(s, i0, j0, k1)], c101 = prev[AT(s, i1, j0, k1)];
                float c011 = prev[AT(s, i0, j1, k1)], c111 = prev[AT(s, i1, j1, k1)];
                float c00 = c000 * (1.0f - sx) + c100 * sx;                            /* :412-415 */
                float c01 = c001 * (1.0f - sx) + c101 * sx;
                float c10 = c010 * (1.0f - sx) + c110 * sx;
                float c11 = c011 * (1.0f - sx) + c111 * sx;
                float c0 = c00 * (1.0f - sy) + c10 * sy;                               /* :417-418 */
                float c1 = c01 * (1.0f - sy) + c11 * sy;
                f[c] = c0 * (1.0f - sz) + c1 * sz;                                     /* :420 */
            }
    fo_set_bounds(s, b, fid);                                                          /* :423 */
}

void fo_inflow_density(fo_sim* s)                                                      /* :65-67 */
{
    for (int j = 1; j <= s->height; ++j)
        for (int k = 1; k <= s->depth; ++k)
            fo_add_density(s, 1, j, k, 0.001f);
}

void fo_inflow_velocity(fo_sim* s)                                                     /* :103-105 */
{
    for (int j = 1; j <= s->height; ++j)
        for (int k = 1; k <= s->depth; ++k)
            fo_set_velocity(s, 1, j, k, (float)s->speed, 0.0f, 0.0f);
}

/* ---- step, simulation.cpp:96-150 (frame write :140-148 left to the caller) */
void fo_step(fo_sim* s)
{
    size_t bytes = (size_t)s->size * sizeof(float);
    fo_inflow_velocity(s);                                                             /* :103-105 */
    memcpy(s->f[FO_VX_PREV], s->f[FO_VX], bytes);                                      /* :108-110 */
    memcpy(s->f[FO_VY_PREV], s->f[FO_VY], bytes);
    memcpy(s->f[FO_VZ_PREV], s->f[FO_VZ], bytes);
    fo_diffuse(s, 1, FO_VX, FO_VX_PREV);                                               /* :115-117 */
    fo_diffuse(s, 2, FO_VY, FO_VY_PREV);
    fo_diffuse(s, 3, FO_VZ, FO_VZ_PREV);
    fo_project(s);                                                                     /* :120 */
    fo_advect(s, 1, FO_VX, FO_VX_PREV);                                                /* :125-127 */
    fo_advect(s, 2, FO_VY, FO_VY_PREV);
    fo_advect(s, 3, FO_VZ, FO_VZ_PREV);
    fo_project(s);                                                                     /* :130 */
    if (!s->elide_dead) fo_diffuse(s, 0, FO_DENS, FO_BUFFER);                          /* :135 (dead, F2) */
    fo_advect(s, 0, FO_DENS, FO_BUFFER);                                               /* :136 */
}

void fo_run_iteration(fo_sim* s)                                                       /* :65-71 */
{
    fo_inflow_density(s);
    memcpy(s->f[FO_BUFFER], s->f[FO_DENS], (size_t)s->size * sizeof(float));           /* :70 */
    fo_step(s);
}

void fo_run_iterations(fo_sim* s, int n) { for (int i = 0; i < n; ++i) fo_run_iteration(s); }
