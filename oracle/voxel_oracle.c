/* voxel_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into or called by the product).
 *
 * Plain-C restatement of the reference's STL voxeliser, object_loader.cpp:288-449
 * (loadSTLIntoObstacles), as the checker for the CUDA voxeliser (csrc/voxelize.cuh).
 *
 * What is restated exactly (same float expressions, same operand order, no contraction):
 *   - the rotation R_x R_y R_z about objCenter          object_loader.cpp:182-202, 311-327
 *   - objCenter: the reference never updates orig_min / orig_max from the mesh, so the centre is
 *     ((1e6f + -1e6f) * 0.5f) = 0 on every axis          object_loader.cpp:301-309  (quirk kept)
 *   - bounding sphere of the UNROTATED vertices, 5 % padding, lattice resolution
 *     max(objSize / 200, 0.02), nx/ny/nz                  object_loader.cpp:329-378
 *   - the coarse 64^3 occupancy grid at 5 x resolution    object_loader.cpp:29-86, 385-393
 *   - Moeller-Trumbore with EPS 1e-6 and t > 1e-3, odd-parity inside test
 *                                                         object_loader.cpp:205-244
 *   - the map from lattice point to simulation cell and the bounds test
 *                                                         object_loader.cpp:425-437
 * What cannot be restated: the reference draws the per-sample jitter and ray direction from a
 * std::minstd_rand seeded with a hash of the OpenMP thread id (object_loader.cpp:399-401), so its mask
 * differs from run to run.  Here both come from a counter-based hash of the sample index (mode 0), or
 * are switched off (mode 1: no jitter, ray (1, 0.1, 0.2) = the first ray of isInsideMesh, :250), or the
 * lattice is replaced by one sample per simulation cell (mode 2, SURVEY 8(f)-2: no porosity).
 * Parity against the reference itself is therefore statistical (tests/test_voxelize.py), parity of the
 * CUDA kernels against this file is bit-exact.
 */
#include <math.h>
#ifndef M_PI
#define M_PI 3.14159265358979323846   /* -std=c11 hides the glibc macro; same double as <cmath>'s */
#endif
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define VO_COARSE 64

typedef struct { float x, y, z; } vo_pt;

/* the lattice / mapping parameters both sides derive from the mesh; exported for the tests */
typedef struct {
    float center[3];
    float max_radius;
    float pmin[3], pmax[3];
    float obj_size, resolution, coarse_res, grid_scale;
    int nx, ny, nz;
} vo_setup;

uint32_t vo_hash(uint32_t idx, uint32_t stream)
{
    uint32_t x = idx * 4u + stream;
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

static vo_pt vo_rotate(vo_pt p, const float m[9])
{
    vo_pt o;
    o.x = m[0] * p.x + m[1] * p.y + m[2] * p.z;      /* object_loader.cpp:197-199 */
    o.y = m[3] * p.x + m[4] * p.y + m[5] * p.z;
    o.z = m[6] * p.x + m[7] * p.y + m[8] * p.z;
    return o;
}

/* object_loader.cpp:183-199: the nine coefficients as the reference forms them */
void vo_rotation(float rx_deg, float ry_deg, float rz_deg, float m[9])
{
    /* deg2rad (:176-178): float * M_PI is evaluated in double, the result narrowed to float */
    float rx = (float)(rx_deg * M_PI / 180.0f), ry = (float)(ry_deg * M_PI / 180.0f), rz = (float)(rz_deg * M_PI / 180.0f);
    float cx = cosf(rx), sx = sinf(rx), cy = cosf(ry), sy = sinf(ry), cz = cosf(rz), sz = sinf(rz);
    m[0] = cy * cz;                 m[1] = -cy * sz;                 m[2] = sy;
    m[3] = sx * sy * cz + cx * sz;  m[4] = -sx * sy * sz + cx * cz;  m[5] = -sx * cy;
    m[6] = -cx * sy * cz + sx * sz; m[7] = cx * sy * sz + sx * cz;   m[8] = cx * cy;
}

static float vo_max3(float a, float b, float c) { float m = a > b ? a : b; return m > c ? m : c; }
static float vo_min3(float a, float b, float c) { float m = a < b ? a : b; return m < c ? m : c; }

/* object_loader.cpp:205-235 */
static int vo_hit(vo_pt o, vo_pt d, const float* t)
{
    const float EPS = 1e-6f;
    vo_pt e1 = { t[3] - t[0], t[4] - t[1], t[5] - t[2] };
    vo_pt e2 = { t[6] - t[0], t[7] - t[1], t[8] - t[2] };
    vo_pt h = { d.y * e2.z - d.z * e2.y, d.z * e2.x - d.x * e2.z, d.x * e2.y - d.y * e2.x };
    float a = e1.x * h.x + e1.y * h.y + e1.z * h.z;
    if (fabsf(a) < EPS) return 0;
    float f = 1.0f / a;
    vo_pt s = { o.x - t[0], o.y - t[1], o.z - t[2] };
    float u = f * (s.x * h.x + s.y * h.y + s.z * h.z);
    if (u < 0.0f || u > 1.0f) return 0;
    vo_pt q = { s.y * e1.z - s.z * e1.y, s.z * e1.x - s.x * e1.z, s.x * e1.y - s.y * e1.x };
    float v = f * (d.x * q.x + d.y * q.y + d.z * q.z);
    if (v < 0.0f || u + v > 1.0f) return 0;
    float tt = f * (e2.x * q.x + e2.y * q.y + e2.z * q.z);
    return tt > 1e-3f;
}

/* object_loader.cpp:297-378 + 425-428 */
void vo_make_setup(const float* tris, int n, int W, int H, int D, float scale, vo_setup* s)
{
    const float omin = 1e6f, omax = -1e6f;                       /* never updated: :297-309 */
    s->center[0] = s->center[1] = s->center[2] = (omin + omax) * 0.5f;
    float m2 = 0.0f;
    for (int i = 0; i < 3 * n; ++i) {                            /* :341-347 */
        float dx = tris[3 * i] - s->center[0], dy = tris[3 * i + 1] - s->center[1], dz = tris[3 * i + 2] - s->center[2];
        float d = dx * dx + dy * dy + dz * dz;
        if (d > m2) m2 = d;
    }
    s->max_radius = sqrtf(m2);
    float pad = s->max_radius * 0.05f;
    for (int a = 0; a < 3; ++a) {
        s->pmin[a] = (s->center[a] - s->max_radius) - pad;
        s->pmax[a] = (s->center[a] + s->max_radius) + pad;
    }
    s->obj_size = vo_max3(s->pmax[0] - s->pmin[0], s->pmax[1] - s->pmin[1], s->pmax[2] - s->pmin[2]);
    float r = s->obj_size / 200.0f;
    s->resolution = r > 0.02f ? r : 0.02f;
    s->nx = (int)((s->pmax[0] - s->pmin[0]) / s->resolution);
    s->ny = (int)((s->pmax[1] - s->pmin[1]) / s->resolution);
    s->nz = (int)((s->pmax[2] - s->pmin[2]) / s->resolution);
    s->coarse_res = s->resolution * 5.0f;
    s->grid_scale = scale * vo_min3((float)W, (float)H, (float)D) / s->obj_size;
}

static int vo_coarse_contains(const unsigned char* occ, const vo_setup* s, float x, float y, float z)
{
    if (x < s->pmin[0] || y < s->pmin[1] || z < s->pmin[2]) return 0;        /* :75-84 */
    int ix = (int)((x - s->pmin[0]) / s->coarse_res);
    int iy = (int)((y - s->pmin[1]) / s->coarse_res);
    int iz = (int)((z - s->pmin[2]) / s->coarse_res);
    if (ix < 0 || ix >= VO_COARSE || iy < 0 || iy >= VO_COARSE || iz < 0 || iz >= VO_COARSE) return 0;
    return occ[ix + iy * VO_COARSE + iz * VO_COARSE * VO_COARSE];
}

/* obs: (D+2)(H+2)(W+2) floats in the reference layout (simulation.h:9), cells set to 1.0f.
 * Returns the reference's `added` count (samples that landed inside the grid, duplicates included),
 * or -1 for an empty mesh. */
long long vo_voxelize(const float* tris, int n, int W, int H, int D, float scale, float rot_x, float rot_y,
                      float rot_z, float tx, float ty, float tz, int mode, float* obs, vo_setup* setup_out)
{
    if (n <= 0) return -1;
    vo_setup s;
    vo_make_setup(tris, n, W, H, D, scale, &s);
    if (setup_out) *setup_out = s;
    float m[9];
    vo_rotation(rot_x, rot_y, rot_z, m);
    float* rt = (float*)malloc((size_t)n * 9 * sizeof(float));
    for (int i = 0; i < 3 * n; ++i) {                            /* :316-326 */
        vo_pt l = { tris[3 * i] - s.center[0], tris[3 * i + 1] - s.center[1], tris[3 * i + 2] - s.center[2] };
        vo_pt r = vo_rotate(l, m);
        rt[3 * i] = r.x + s.center[0]; rt[3 * i + 1] = r.y + s.center[1]; rt[3 * i + 2] = r.z + s.center[2];
    }
    unsigned char* occ = (unsigned char*)calloc((size_t)VO_COARSE * VO_COARSE * VO_COARSE, 1);
    for (int i = 0; i < n; ++i) {                                /* VoxelGrid::insert, :53-73 */
        const float* t = rt + 9 * i;
        float lo[3], hi[3];
        for (int a = 0; a < 3; ++a) { lo[a] = vo_min3(t[a], t[3 + a], t[6 + a]); hi[a] = vo_max3(t[a], t[3 + a], t[6 + a]); }
        int c0[3], c1[3];
        for (int a = 0; a < 3; ++a) {
            int v0 = (int)((lo[a] - s.pmin[a]) / s.coarse_res), v1 = (int)((hi[a] - s.pmin[a]) / s.coarse_res);
            c0[a] = v0 > 0 ? v0 : 0; c1[a] = v1 < VO_COARSE - 1 ? v1 : VO_COARSE - 1;
        }
        for (int iz = c0[2]; iz <= c1[2]; ++iz)
            for (int iy = c0[1]; iy <= c1[1]; ++iy)
                for (int ix = c0[0]; ix <= c1[0]; ++ix) occ[ix + iy * VO_COARSE + iz * VO_COARSE * VO_COARSE] = 1;
    }
    const float gcx = (float)W / 2, gcy = (float)H / 2, gcz = (float)D / 2;     /* :428 */
    long long added = 0;
    const long long total = mode == 2 ? (long long)W * H * D : (long long)s.nx * s.ny * s.nz;
    for (long long idx = 0; idx < total; ++idx) {
        vo_pt p, dir;
        int gx, gy, gz;
        if (mode == 2) {                                         /* one sample per simulation cell: its centre */
            gx = 1 + (int)(idx % W); gy = 1 + (int)((idx / W) % H); gz = 1 + (int)(idx / ((long long)W * H));
            p.x = ((((float)gx + 0.5f) - gcx) - tx) / s.grid_scale + s.center[0];
            p.y = ((((float)gy + 0.5f) - gcy) - ty) / s.grid_scale + s.center[1];
            p.z = ((((float)gz + 0.5f) - gcz) - tz) / s.grid_scale + s.center[2];
        } else {                                                 /* lattice point (i, j, k), k fastest: :405-412 */
            int k = (int)(idx % s.nz), j = (int)((idx / s.nz) % s.ny), i = (int)(idx / ((long long)s.nz * s.ny));
            p.x = s.pmin[0] + (float)i * s.resolution;
            p.y = s.pmin[1] + (float)j * s.resolution;
            p.z = s.pmin[2] + (float)k * s.resolution;
        }
        if (!vo_coarse_contains(occ, &s, p.x, p.y, p.z)) continue;            /* :414-417 */
        if (mode == 0) {                                         /* :419-425, hash instead of minstd_rand */
            p.x += (float)(vo_hash((uint32_t)idx, 0) % 1000u) * 1e-6f - 5e-4f;
            p.y += (float)(vo_hash((uint32_t)idx, 1) % 1000u) * 1e-6f - 5e-4f;
            p.z += (float)(vo_hash((uint32_t)idx, 2) % 1000u) * 1e-6f - 5e-4f;
            uint32_t r = vo_hash((uint32_t)idx, 3);
            dir.x = 0.1f + 0.9f * ((float)(r & 1023u) * (1.0f / 1024.0f));
            dir.y = 0.1f + 0.9f * ((float)((r >> 10) & 1023u) * (1.0f / 1024.0f));
            dir.z = 0.1f + 0.9f * ((float)((r >> 20) & 1023u) * (1.0f / 1024.0f));
        } else {
            dir.x = 1.0f; dir.y = 0.1f; dir.z = 0.2f;
        }
        int hits = 0;
        for (int t = 0; t < n; ++t) hits += vo_hit(p, dir, rt + 9 * t);        /* :238-246 */
        if ((hits & 1) == 0) continue;
        if (mode != 2) {                                         /* :430-432 */
            gx = (int)((p.x - s.center[0]) * s.grid_scale + gcx + tx);
            gy = (int)((p.y - s.center[1]) * s.grid_scale + gcy + ty);
            gz = (int)((p.z - s.center[2]) * s.grid_scale + gcz + tz);
        }
        if (gx >= 1 && gx <= W && gy >= 1 && gy <= H && gz >= 1 && gz <= D) {  /* :434-441 */
            obs[(size_t)gx + (size_t)(W + 2) * ((size_t)gy + (size_t)(H + 2) * (size_t)gz)] = 1.0f;
            ++added;
        }
    }
    free(occ); free(rt);
    return added;
}
