// voxelize.cuh -- the reference's STL voxeliser (object_loader.cpp:288-449, loadSTLIntoObstacles) as
// CUDA kernels: SURVEY 8(f)-2.  The mesh (n x 9 floats) comes from the host STL reader
// (host/object_loader.cpp, fluid_simulation_b200/stl.py); everything from the rotation to the obstacle
// mask runs on the device and the mask is written straight into the pitched `obs` field.
//
//   k_vox_rotate   rotate every vertex about objCenter (R_x R_y R_z, :182-202, 311-327) and reduce the
//                  bounding-sphere radius^2 of the UNROTATED vertices (:341-347; atomicMax on the bits of a
//                  non-negative float is order-independent)
//   k_vox_coarse   VoxelGrid::insert (:53-73): triangle bounding boxes into the 64^3 occupancy grid
//   k_vox_select   lattice points (or simulation cells) that pass VoxelGrid::contains (:75-86, 414-417)
//                  are appended to a candidate list
//   k_vox_inside   one thread per candidate: jitter + ray (:419-425), Moeller-Trumbore against every
//                  triangle (:205-246; triangles staged through shared memory), odd parity -> map to the
//                  simulation cell (:425-441) and set obs = 1 (Simulation::addObstacle, simulation.cpp:155)
//
// Arithmetic keeps the reference's float expressions and operand order (the library is built with
// --fmad=false), so the mask is bit-identical to the plain-C restatement the tests check it against.  The
// reference's own mask is not reproducible (its jitter and ray come from a minstd_rand seeded with the
// OpenMP thread id, :399); the sampling modes below replace that:
//   FLUID_VOX_HASHED  jitter and ray from a counter-based hash of the sample index (same distributions)
//   FLUID_VOX_FIXED   no jitter, ray (1, 0.1, 0.2) (the first ray of isInsideMesh, :250)
//   FLUID_VOX_CELLS   one sample per simulation cell (its centre mapped back into mesh space) instead of
//                     the 200^3 lattice: no porous masks when the grid is finer than the lattice
#pragma once
#include "handle.cuh"
#include <math.h>

#define VOX_COARSE 64
#define VOX_TILE 256                       // triangles per shared-memory tile

struct VoxSetup {                          // object_loader.cpp:297-378, 425-428
    float cx, cy, cz;                      // objCenter
    float pminx, pminy, pminz;
    float resolution, coarse_res, grid_scale;
    float gcx, gcy, gcz;                   // gridCenter
    float tx, ty, tz;
    int nx, ny, nz;
    int mode;
};

__host__ __device__ inline unsigned vox_hash(unsigned idx, unsigned stream)
{
    unsigned x = idx * 4u + stream;
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

__global__ void k_vox_rotate(int nverts, const float* __restrict__ in, float* __restrict__ out,
                             float cx, float cy, float cz, const float* __restrict__ m, unsigned* __restrict__ max_d2)
{
    unsigned best = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nverts; i += gridDim.x * blockDim.x) {
        const float lx = in[3 * i] - cx, ly = in[3 * i + 1] - cy, lz = in[3 * i + 2] - cz;
        out[3 * i]     = (m[0] * lx + m[1] * ly + m[2] * lz) + cx;
        out[3 * i + 1] = (m[3] * lx + m[4] * ly + m[5] * lz) + cy;
        out[3 * i + 2] = (m[6] * lx + m[7] * ly + m[8] * lz) + cz;
        const float d = lx * lx + ly * ly + lz * lz;
        best = max(best, __float_as_uint(d));
    }
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((threadIdx.x & 31) == 0 && best) atomicMax(max_d2, best);
}

__global__ void k_vox_coarse(int n, const float* __restrict__ rt, VoxSetup s, unsigned char* __restrict__ occ)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* t = rt + 9 * (long long)i;
    const float pmin[3] = { s.pminx, s.pminy, s.pminz };
    int c0[3], c1[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = fminf(fminf(t[a], t[3 + a]), t[6 + a]), hi = fmaxf(fmaxf(t[a], t[3 + a]), t[6 + a]);
        c0[a] = max(0, (int)((lo - pmin[a]) / s.coarse_res));
        c1[a] = min(VOX_COARSE - 1, (int)((hi - pmin[a]) / s.coarse_res));
    }
    for (int iz = c0[2]; iz <= c1[2]; ++iz)
        for (int iy = c0[1]; iy <= c1[1]; ++iy)
            for (int ix = c0[0]; ix <= c1[0]; ++ix) occ[ix + iy * VOX_COARSE + iz * VOX_COARSE * VOX_COARSE] = 1;
}

// sample `idx` in mesh space (before the jitter); CELLS mode also yields the simulation cell
__device__ __forceinline__ void vox_sample(const VoxSetup& s, int W, int H, long long idx, float& px, float& py, float& pz,
                                           int& gx, int& gy, int& gz)
{
    if (s.mode == FLUID_VOX_CELLS) {
        gx = 1 + (int)(idx % W); gy = 1 + (int)((idx / W) % H); gz = 1 + (int)(idx / ((long long)W * H));
        px = ((((float)gx + 0.5f) - s.gcx) - s.tx) / s.grid_scale + s.cx;
        py = ((((float)gy + 0.5f) - s.gcy) - s.ty) / s.grid_scale + s.cy;
        pz = ((((float)gz + 0.5f) - s.gcz) - s.tz) / s.grid_scale + s.cz;
    } else {
        const int k = (int)(idx % s.nz), j = (int)((idx / s.nz) % s.ny), i = (int)(idx / ((long long)s.nz * s.ny));
        px = s.pminx + (float)i * s.resolution;
        py = s.pminy + (float)j * s.resolution;
        pz = s.pminz + (float)k * s.resolution;
        gx = gy = gz = 0;
    }
}

__global__ void k_vox_select(VoxSetup s, int W, int H, long long total, const unsigned char* __restrict__ occ,
                             unsigned* __restrict__ cand, unsigned* __restrict__ ncand)
{
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        float px, py, pz; int gx, gy, gz;
        vox_sample(s, W, H, idx, px, py, pz, gx, gy, gz);
        bool keep = !(px < s.pminx || py < s.pminy || pz < s.pminz);
        if (keep) {
            const int ix = (int)((px - s.pminx) / s.coarse_res), iy = (int)((py - s.pminy) / s.coarse_res),
                      iz = (int)((pz - s.pminz) / s.coarse_res);
            keep = ix >= 0 && ix < VOX_COARSE && iy >= 0 && iy < VOX_COARSE && iz >= 0 && iz < VOX_COARSE &&
                   occ[ix + iy * VOX_COARSE + iz * VOX_COARSE * VOX_COARSE];
        }
        // warp-aggregated append (the list order is irrelevant: every candidate is independent)
        const unsigned ballot = __ballot_sync(__activemask(), keep);
        if (keep) {
            const unsigned lane = threadIdx.x & 31u, leader = __ffs(ballot) - 1;
            unsigned base = 0;
            if (lane == leader) base = atomicAdd(ncand, __popc(ballot));
            base = __shfl_sync(ballot, base, leader);
            cand[base + __popc(ballot & ((1u << lane) - 1u))] = (unsigned)idx;
        }
    }
}

__global__ void __launch_bounds__(VOX_TILE) k_vox_inside(VoxSetup s, Grid g, int W, int H, int D, int ntris,
                                                         const float* __restrict__ rt, const unsigned* __restrict__ cand,
                                                         const unsigned* __restrict__ ncand, float* __restrict__ obs,
                                                         unsigned long long* __restrict__ added)
{
    __shared__ float tri[VOX_TILE * 9];
    const unsigned n = *ncand;
    for (unsigned c0 = blockIdx.x * VOX_TILE; c0 < n; c0 += gridDim.x * VOX_TILE) {
        const unsigned c = c0 + threadIdx.x;
        const bool live = c < n;
        const long long idx = live ? (long long)cand[c] : 0;
        float px, py, pz, dx, dy, dz; int gx, gy, gz;
        vox_sample(s, W, H, idx, px, py, pz, gx, gy, gz);
        if (s.mode == FLUID_VOX_HASHED) {                         // :419-425
            px += (float)(vox_hash((unsigned)idx, 0) % 1000u) * 1e-6f - 5e-4f;
            py += (float)(vox_hash((unsigned)idx, 1) % 1000u) * 1e-6f - 5e-4f;
            pz += (float)(vox_hash((unsigned)idx, 2) % 1000u) * 1e-6f - 5e-4f;
            const unsigned r = vox_hash((unsigned)idx, 3);
            dx = 0.1f + 0.9f * ((float)(r & 1023u) * (1.0f / 1024.0f));
            dy = 0.1f + 0.9f * ((float)((r >> 10) & 1023u) * (1.0f / 1024.0f));
            dz = 0.1f + 0.9f * ((float)((r >> 20) & 1023u) * (1.0f / 1024.0f));
        } else { dx = 1.0f; dy = 0.1f; dz = 0.2f; }
        int hits = 0;
        for (int t0 = 0; t0 < ntris; t0 += VOX_TILE) {
            const int nt = min(VOX_TILE, ntris - t0);
            __syncthreads();
            for (int i = threadIdx.x; i < nt * 9; i += VOX_TILE) tri[i] = rt[(long long)t0 * 9 + i];
            __syncthreads();
            if (!live) continue;
            for (int t = 0; t < nt; ++t) {                        // rayIntersectsTriangle, :205-235
                const float* v = tri + 9 * t;
                const float e1x = v[3] - v[0], e1y = v[4] - v[1], e1z = v[5] - v[2];
                const float e2x = v[6] - v[0], e2y = v[7] - v[1], e2z = v[8] - v[2];
                const float hx = dy * e2z - dz * e2y, hy = dz * e2x - dx * e2z, hz = dx * e2y - dy * e2x;
                const float a = e1x * hx + e1y * hy + e1z * hz;
                if (fabsf(a) < 1e-6f) continue;
                const float f = 1.0f / a;
                const float sx = px - v[0], sy = py - v[1], sz = pz - v[2];
                const float u = f * (sx * hx + sy * hy + sz * hz);
                if (u < 0.0f || u > 1.0f) continue;
                const float qx = sy * e1z - sz * e1y, qy = sz * e1x - sx * e1z, qz = sx * e1y - sy * e1x;
                const float w = f * (dx * qx + dy * qy + dz * qz);
                if (w < 0.0f || u + w > 1.0f) continue;
                const float tt = f * (e2x * qx + e2y * qy + e2z * qz);
                hits += tt > 1e-3f;
            }
        }
        bool in = live && (hits & 1);
        if (in && s.mode != FLUID_VOX_CELLS) {                    // :425-432
            gx = (int)((px - s.cx) * s.grid_scale + s.gcx + s.tx);
            gy = (int)((py - s.cy) * s.grid_scale + s.gcy + s.ty);
            gz = (int)((pz - s.cz) * s.grid_scale + s.gcz + s.tz);
        }
        in = in && gx >= 1 && gx <= W && gy >= 1 && gy <= H && gz >= 1 && gz <= D;
        if (in) {
            const int zl = gz - g.zlo + 1;                        // z-slabs: every rank keeps the planes it holds
            if (zl >= g.zl_lo && zl <= g.zl_hi) obs[fl_idx(g, gx, gy, zl)] = 1.0f;
        }
        const unsigned cnt = __popc(__ballot_sync(0xffffffffu, in));
        if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(added, (unsigned long long)cnt);
    }
}

// deg2rad + the nine coefficients of R_x R_y R_z as object_loader.cpp:176-199 forms them (host floats)
static void vox_rotation(float rx_deg, float ry_deg, float rz_deg, float m[9])
{
    const float rx = (float)(rx_deg * M_PI / 180.0f), ry = (float)(ry_deg * M_PI / 180.0f), rz = (float)(rz_deg * M_PI / 180.0f);
    const float cx = cosf(rx), sx = sinf(rx), cy = cosf(ry), sy = sinf(ry), cz = cosf(rz), sz = sinf(rz);
    m[0] = cy * cz;                 m[1] = -cy * sz;                 m[2] = sy;
    m[3] = sx * sy * cz + cx * sz;  m[4] = -sx * sy * sz + cx * cz;  m[5] = -sx * cy;
    m[6] = -cx * sy * cz + sx * sz; m[7] = cx * sy * sz + sx * cz;   m[8] = cx * cy;
}

// loadSTLIntoObstacles from the triangle list on: obs cells of voxels inside the mesh := 1
int fl_voxelize(fluid_handle* h, const float* tris, int ntris, float scale, float rot_x, float rot_y, float rot_z,
                float tx, float ty, float tz, int mode, long long* added_out)
{
    if (added_out) *added_out = 0;
    if (!tris || ntris <= 0) return fl_fail(h, FLUID_E_ARG, "fluid_voxelize: empty mesh (object_loader.cpp:292-296)");
    if (mode != FLUID_VOX_HASHED && mode != FLUID_VOX_FIXED && mode != FLUID_VOX_CELLS)
        return fl_fail(h, FLUID_E_ARG, "fluid_voxelize: unknown sampling mode %d", mode);
    const Grid& g = h->g;
    const int nverts = 3 * ntris;
    float m[9];
    vox_rotation(rot_x, rot_y, rot_z, m);
    float *d_in = nullptr, *d_rt = nullptr, *d_m = nullptr;
    unsigned char* d_occ = nullptr;
    unsigned *d_u = nullptr, *d_cand = nullptr;                     // d_u: [0] max d^2 bits, [1] candidate count, [2..3] added
    auto release = [&]() { cudaFree(d_in); cudaFree(d_rt); cudaFree(d_m); cudaFree(d_occ); cudaFree(d_u); cudaFree(d_cand); };
#define VOX_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { release();                      \
        return fl_fail(h, FLUID_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } } while (0)
    VOX_CUDA(cudaMalloc(&d_in, (size_t)nverts * 3 * sizeof(float)));
    VOX_CUDA(cudaMalloc(&d_rt, (size_t)nverts * 3 * sizeof(float)));
    VOX_CUDA(cudaMalloc(&d_m, 9 * sizeof(float)));
    VOX_CUDA(cudaMalloc(&d_occ, (size_t)VOX_COARSE * VOX_COARSE * VOX_COARSE));
    VOX_CUDA(cudaMalloc(&d_u, 4 * sizeof(unsigned)));
    VOX_CUDA(cudaMemcpyAsync(d_in, tris, (size_t)nverts * 3 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    VOX_CUDA(cudaMemcpyAsync(d_m, m, sizeof m, cudaMemcpyHostToDevice, h->stream));
    VOX_CUDA(cudaMemsetAsync(d_occ, 0, (size_t)VOX_COARSE * VOX_COARSE * VOX_COARSE, h->stream));
    VOX_CUDA(cudaMemsetAsync(d_u, 0, 4 * sizeof(unsigned), h->stream));

    VoxSetup s = {};
    s.mode = mode;
    const float omin = 1e6f, omax = -1e6f;                          // never updated by the reference (:297-309)
    s.cx = s.cy = s.cz = (omin + omax) * 0.5f;
    k_vox_rotate<<<fl_grid_1d(nverts), 256, 0, h->stream>>>(nverts, d_in, d_rt, s.cx, s.cy, s.cz, d_m, d_u);
    h->launches++;
    unsigned bits = 0;
    VOX_CUDA(cudaMemcpyAsync(&bits, d_u, sizeof bits, cudaMemcpyDeviceToHost, h->stream));
    VOX_CUDA(cudaStreamSynchronize(h->stream));
    float d2; memcpy(&d2, &bits, sizeof d2);
    // :349-378: bounding cube of the sphere, 5 % padding, lattice resolution
    const float max_radius = sqrtf(d2), pad = max_radius * 0.05f;
    const float pmin[3] = { (s.cx - max_radius) - pad, (s.cy - max_radius) - pad, (s.cz - max_radius) - pad };
    const float pmax[3] = { (s.cx + max_radius) + pad, (s.cy + max_radius) + pad, (s.cz + max_radius) + pad };
    const float obj_size = fmaxf(fmaxf(pmax[0] - pmin[0], pmax[1] - pmin[1]), pmax[2] - pmin[2]);
    if (!(obj_size > 0.0f)) { release(); return fl_fail(h, FLUID_E_ARG, "fluid_voxelize: degenerate mesh (all vertices at the centre)"); }
    const float r = obj_size / 200.0f;
    s.pminx = pmin[0]; s.pminy = pmin[1]; s.pminz = pmin[2];
    s.resolution = r > 0.02f ? r : 0.02f;
    s.nx = (int)((pmax[0] - pmin[0]) / s.resolution);
    s.ny = (int)((pmax[1] - pmin[1]) / s.resolution);
    s.nz = (int)((pmax[2] - pmin[2]) / s.resolution);
    s.coarse_res = s.resolution * 5.0f;
    s.grid_scale = scale * fminf(fminf((float)g.W, (float)g.H), (float)g.D) / obj_size;     // :426-427
    s.gcx = (float)g.W / 2; s.gcy = (float)g.H / 2; s.gcz = (float)g.D / 2;                 // :428
    s.tx = tx; s.ty = ty; s.tz = tz;
    const long long total = mode == FLUID_VOX_CELLS ? (long long)g.W * g.H * g.D : (long long)s.nx * s.ny * s.nz;
    if (total <= 0 || total > 0xffffffffLL) { release(); return fl_fail(h, FLUID_E_ARG, "fluid_voxelize: %lld samples out of range", total); }
    VOX_CUDA(cudaMalloc(&d_cand, (size_t)total * sizeof(unsigned)));

    k_vox_coarse<<<(ntris + 127) / 128, 128, 0, h->stream>>>(ntris, d_rt, s, d_occ);
    k_vox_select<<<fl_grid_1d(total, 256, 148 * 32), 256, 0, h->stream>>>(s, g.W, g.H, total, d_occ, d_cand, d_u + 1);
    k_vox_inside<<<148 * 8, VOX_TILE, 0, h->stream>>>(s, g, g.W, g.H, g.D, ntris, d_rt, d_cand, d_u + 1,
                                                      h->field[FLUID_OBS], (unsigned long long*)(d_u + 2));
    h->launches += 3;
    unsigned long long added = 0;
    VOX_CUDA(cudaMemcpyAsync(&added, d_u + 2, sizeof added, cudaMemcpyDeviceToHost, h->stream));
    VOX_CUDA(cudaStreamSynchronize(h->stream));
    VOX_CUDA(cudaGetLastError());
#undef VOX_CUDA
    release();
    h->flags_dirty = true;
    if (added_out) *added_out = (long long)added;
    return 0;
}
