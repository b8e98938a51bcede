// tiled.cuh -- TMA-staged shared-memory sweep kernel: one full linearSolver sweep
// (simulation.cpp:259-271: relax every interior cell, then setBounds) per pass over HBM.
//
// 2.5-D blocking: a CTA owns an x/y tile of TQ x-quads (128 cells) by TY rows and streams along z
// through a chunk of planes.  Each plane tile (+1 quad halo in x, +2 rows in y) is brought into a
// ring of NSLOT shared-memory slots by ONE 3-D TMA box load (cp.async.bulk.tensor, zero fill
// outside the array) signalled on an mbarrier, prefetched one plane ahead of the compute.
//   red-black: iteration p relaxes the colour-0 cells of plane p-1 IN PLACE in shared memory
//     (tile + 1 halo ring, from planes p-2..p), then relaxes the colour-1 cells of plane p-2 from
//     the colour-0 values of planes p-3..p-1 and hands the finished quad to the fused setBounds
//     epilogue (fl_store_bounded) -- both half sweeps and the bounds cost one read of x, one read
//     of the right-hand side and one write: the 12 B/cell/sweep algorithmic minimum.
//   Jacobi: same pipeline without the in-place stage.
// Solid cells are relaxed like any other cell and only zeroed by the epilogue, so colour-1 cells
// see the transient colour-0 value of a solid neighbour exactly as in the ordering-matched CPU
// loop.  Arithmetic is fl_relax (no FMA, reference operand order): bit-identical to the simple
// kernels and to the CPU oracle.
#pragma once
#include "handle.cuh"
#include "kernels_stencil.cuh"
#include <cuda.h>
#include <map>
#include <cstdlib>

#define TL_TQ     32                     // owned x-quads per tile row (one warp wide)
#define TL_TY     30                     // owned rows per tile
#define TL_BOX_X  (4 * (TL_TQ + 2))      // 136 floats: halo quad + 32 quads + halo quad
#define TL_BOX_Y  (TL_TY + 4)            // 34 rows: 2 halo rows each side
#define TL_NSLOT  5
#define TL_SLOT_FLOATS 4640              // 136*34 = 4624 floats, padded to a multiple of 128 B
#define TL_BOX_BYTES (TL_BOX_X * TL_BOX_Y * 4)
#define TL_THREADS 256

struct alignas(64) TiledArgs {
    CUtensorMap map[3];                  // source field of each batched solve
    const float* rhs[3];
    float* dst[3];
    int b[3];
    const unsigned char* flags;
    Grid g;
    float a, c_recip;
    int ntx, zc;
};

// ---- PTX wrappers (mbarrier + TMA) -----------------------------------------------------------
__device__ __forceinline__ unsigned tl_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tl_mbar_init(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tl_smem(bar)), "r"(count));
}
__device__ __forceinline__ void tl_mbar_expect_tx(unsigned long long* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tl_smem(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tl_mbar_wait(unsigned long long* bar, unsigned parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n"
        "W_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra W_%=;\n\t}"
        ::"r"(tl_smem(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tl_tma_load_3d(void* dst, const CUtensorMap* map, int c0, int c1, int c2,
                                               unsigned long long* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(tl_smem(dst)), "l"(reinterpret_cast<unsigned long long>(map)), "r"(c0), "r"(c1), "r"(c2),
          "r"(tl_smem(bar)) : "memory");
}
__device__ __forceinline__ void tl_fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ float4 tl_lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }

// the 7-point neighbourhood of one x-quad in shared memory
struct TlQuad { float4 c, yp, ym, zp, zm; float l, r; };

__device__ __forceinline__ void tl_load_quad(const float* __restrict__ zm, const float* __restrict__ zc,
                                             const float* __restrict__ zp, int off, TlQuad& q)
{
    q.c = tl_lds4(zc + off);
    q.l = zc[off - 1];
    q.r = zc[off + 4];
    q.yp = tl_lds4(zc + off + TL_BOX_X);
    q.ym = tl_lds4(zc + off - TL_BOX_X);
    q.zp = tl_lds4(zp + off);
    q.zm = tl_lds4(zm + off);
}

// relax elements (e0, e0+2) of the quad, e0 in {0,1}: the two cells of one colour
__device__ __forceinline__ void tl_relax_pair(const TlQuad& q, const float4& r, float a, float cr, int e0,
                                              float& v0, float& v1)
{
    if (e0 == 0) {
        v0 = fl_relax(r.x, a, cr, q.c.y, q.l,   q.yp.x, q.ym.x, q.zp.x, q.zm.x);
        v1 = fl_relax(r.z, a, cr, q.c.w, q.c.y, q.yp.z, q.ym.z, q.zp.z, q.zm.z);
    } else {
        v0 = fl_relax(r.y, a, cr, q.c.z, q.c.x, q.yp.y, q.ym.y, q.zp.y, q.zm.y);
        v1 = fl_relax(r.w, a, cr, q.r,   q.c.z, q.yp.w, q.ym.w, q.zp.w, q.zm.w);
    }
}

// MODE 0 = Jacobi, 1 = red-black (colour 0 = (x+y+z) even first)
template <int MODE>
__global__ void __launch_bounds__(TL_THREADS, 2) k_sweep_tiled(const __grid_constant__ TiledArgs A)
{
    extern __shared__ __align__(128) float tl_ring[];
    __shared__ unsigned long long bars[TL_NSLOT];

    const Grid& g = A.g;
    const int bx = blockIdx.x % A.ntx, by = blockIdx.x / A.ntx;
    const int batch = blockIdx.z;
    const int za = 1 + blockIdx.y * A.zc;
    const int zb = min(za + A.zc - 1, g.Dl);
    const int cq0 = 1 + bx * TL_TQ;                 // first owned column-quad (cols 4*cq .. 4*cq+3)
    const int y0 = 1 + by * TL_TY;                  // first owned row
    constexpr int HZ = (MODE == 1) ? 2 : 1;
    const int p_first = max(za - HZ, 0), p_last = min(zb + HZ, g.Dl + 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float a = A.a, cr = A.c_recip;
    const float* __restrict__ rhs = A.rhs[batch];
    const int b = A.b[batch];

    if (tid == 0) {
        for (int i = 0; i < TL_NSLOT; ++i) tl_mbar_init(&bars[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto slot_of = [&](int p) { return (p - p_first) % TL_NSLOT; };
    auto plane = [&](int p) -> float* { return tl_ring + slot_of(p) * TL_SLOT_FLOATS; };
    auto issue = [&](int p) {
        unsigned long long* bar = &bars[slot_of(p)];
        tl_mbar_expect_tx(bar, TL_BOX_BYTES);
        tl_tma_load_3d(plane(p), &A.map[batch], 4 * (cq0 - 1), y0 - 2, p, bar);
    };
    if (tid == 0) issue(p_first);

    const int xg0 = 4 * (cq0 + lane) - 3;           // first cell of this lane's quad (always odd)
    const int soff = 4 * (lane + 1);                // its column inside a shared-memory row
    const bool lane_in = xg0 <= g.W;

    // planes p_first..p_last are loaded; the pipeline drains until p = zb + HZ (the last chunk has no
    // plane beyond the ghost plane Dl+1 to load, but its final stage still lags HZ iterations behind)
    for (int p = p_first; p <= zb + HZ; ++p) {
        if (tid == 0 && p + 1 <= p_last) issue(p + 1);          // prefetch one plane ahead
        if (p <= p_last) tl_mbar_wait(&bars[slot_of(p)], ((p - p_first) / TL_NSLOT) & 1);

        if (MODE == 1) {
            // ---- colour-0 half sweep of plane zr, in place in shared memory -------------------
            const int zr = p - 1;
            if (zr >= max(za - 1, 1) && zr <= min(zb + 1, g.Dl)) {
                float* pc = plane(zr);
                const float* pm = plane(zr - 1);
                const float* pp = plane(zr + 1);
                const int gz = g.zlo + zr - 1;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int sr = 1 + warp + 8 * k;            // shared-memory rows 1..32
                    const int y = y0 - 2 + sr;
                    if (y < 1 || y > g.H) continue;             // warp-uniform
                    const int e0 = (1 + y + gz) & 1;            // colour of element 0; colour-0 pair
                    const int off = sr * TL_BOX_X + soff;
                    if (lane_in) {
                        TlQuad q; tl_load_quad(pm, pc, pp, off, q);
                        const float4 r = ld4(rhs + fl_idx(g, xg0, y, zr));
                        float v0, v1;
                        tl_relax_pair(q, r, a, cr, e0, v0, v1);
                        // cells beyond W inside a ragged last quad stay untouched (never read)
                        if (e0 == 0) { q.c.x = v0; if (xg0 + 2 <= g.W) q.c.z = v1; }
                        else { if (xg0 + 1 <= g.W) q.c.y = v0; if (xg0 + 3 <= g.W) q.c.w = v1; }
                        *reinterpret_cast<float4*>(pc + off) = q.c;
                    }
                    // halo ring in x: the single cell next to the tile on each side
                    if (lane == 0 || lane == 31) {
                        const int xe = (lane == 0) ? 4 * cq0 - 4 : 4 * (cq0 + TL_TQ) - 3;
                        const int ce = (lane == 0) ? 3 : 4 * (TL_TQ + 1);
                        if (xe >= 1 && xe <= g.W && ((xe + y + gz) & 1) == 0) {
                            const int o = sr * TL_BOX_X + ce;
                            pc[o] = fl_relax(rhs[fl_idx(g, xe, y, zr)], a, cr, pc[o + 1], pc[o - 1],
                                             pc[o + TL_BOX_X], pc[o - TL_BOX_X], pp[o], pm[o]);
                        }
                    }
                }
            }
            __syncthreads();
        }

        // ---- final stage of plane zo: colour-1 half sweep (or the whole Jacobi sweep) + setBounds
        const int zo = p - HZ;
        if (zo >= za && zo <= zb) {
            const float* pc = plane(zo);
            const float* pm = plane(zo - 1);
            const float* pp = plane(zo + 1);
            const int gz = g.zlo + zo - 1;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int sr = 2 + warp + 8 * k;                // shared-memory rows 2..31
                const int y = y0 - 2 + sr;
                if (sr > TL_TY + 1 || y > g.H) continue;        // warp-uniform
                if (!lane_in) continue;
                const int off = sr * TL_BOX_X + soff;
                TlQuad q; tl_load_quad(pm, pc, pp, off, q);
                const float4 r = ld4(rhs + fl_idx(g, xg0, y, zo));
                float raw[4];
                if (MODE == 1) {
                    const int e1 = (y + gz) & 1;                // colour-1 pair: 1 - colour of element 0
                    float v0, v1;
                    tl_relax_pair(q, r, a, cr, e1, v0, v1);
                    if (e1 == 0) { raw[0] = v0; raw[1] = q.c.y; raw[2] = v1; raw[3] = q.c.w; }
                    else         { raw[0] = q.c.x; raw[1] = v0; raw[2] = q.c.z; raw[3] = v1; }
                } else {
                    float t0, t1, t2, t3;
                    tl_relax_pair(q, r, a, cr, 0, t0, t2);
                    tl_relax_pair(q, r, a, cr, 1, t1, t3);
                    raw[0] = t0; raw[1] = t1; raw[2] = t2; raw[3] = t3;
                }
                fl_store_bounded(g, A.dst[batch], A.flags, b, xg0, y, zo, raw);
            }
        }
        tl_fence_proxy_async();        // generic-proxy smem accesses before the next TMA overwrite
        __syncthreads();
    }
}

// ---- host side -------------------------------------------------------------------------------
typedef CUresult (*tl_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                 CUtensorMapFloatOOBfill);

struct TiledState {
    tl_encode_fn encode = nullptr;
    std::map<const float*, CUtensorMap> maps;      // one 3-D tensor map per resident field buffer
    bool attr_set = false;
    int zc_override = 0;
};

inline TiledState* tl_state(fluid_handle* h)
{
    if (!h->tiled) {
        TiledState* s = new TiledState();
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            s->encode = (tl_encode_fn)fn;
        if (const char* e = getenv("FLUID_TILED_ZC")) s->zc_override = atoi(e);
        h->tiled = s;
    }
    return (TiledState*)h->tiled;
}

inline void fl_tiled_release(fluid_handle* h)
{
    delete (TiledState*)h->tiled;
    h->tiled = nullptr;
}

inline int tl_get_map(fluid_handle* h, const float* buf, CUtensorMap* out)
{
    TiledState* s = tl_state(h);
    auto it = s->maps.find(buf);
    if (it == s->maps.end()) {
        const Grid& g = h->g;
        CUtensorMap m;
        cuuint64_t dims[3] = { (cuuint64_t)g.pitch, (cuuint64_t)g.rows, (cuuint64_t)(g.Dl + 2) };
        cuuint64_t strides[2] = { (cuuint64_t)g.pitch * 4, (cuuint64_t)g.plane * 4 };
        cuuint32_t box[3] = { TL_BOX_X, TL_BOX_Y, 1 };
        cuuint32_t estr[3] = { 1, 1, 1 };
        CUresult r = s->encode(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)buf, dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fl_fail(h, FLUID_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
        it = s->maps.emplace(buf, m).first;
    }
    *out = it->second;
    return 0;
}

// linearSolver through the tiled kernel; *handled = false leaves the solve to the simple kernels
inline int fl_tiled_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                          float a, float c, bool* handled)
{
    *handled = false;
    const Grid& g = h->g;
    const unsigned solver = h->flags & FLUID_SOLVER_MASK;
    if (solver != FLUID_SOLVER_RB && solver != FLUID_SOLVER_JACOBI) return 0;
    if (h->nranks > 1) return 0;                         // slab mode exchanges halos per sweep (simple path)
    if (!(h->flags & FLUID_TILED_ALWAYS) && (g.W < 96 || g.H < 8 || g.Dl < 4)) return 0;
    TiledState* s = tl_state(h);
    if (!s->encode) return fl_fail(h, FLUID_E_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
    const size_t smem = (size_t)TL_NSLOT * TL_SLOT_FLOATS * sizeof(float);
    if (!s->attr_set) {
        FL_CUDA(h, cudaFuncSetAttribute(k_sweep_tiled<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FL_CUDA(h, cudaFuncSetAttribute(k_sweep_tiled<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        s->attr_set = true;
    }
    TiledArgs A;
    A.g = g; A.a = a; A.c_recip = 1.0f / c; A.flags = h->cflags;         // simulation.cpp:257
    A.ntx = (g.nxq + TL_TQ - 1) / TL_TQ;
    const int nty = (g.H + TL_TY - 1) / TL_TY;
    // z-chunks: enough CTAs for one (single field) or two (batched) full waves of 2 CTAs per SM
    const int slots = 2 * h->sm_count;
    const int target = (nb == 1) ? slots : 2 * slots;
    int nchunks = (target + A.ntx * nty * nb - 1) / (A.ntx * nty * nb);
    if (nchunks < 1) nchunks = 1;
    int zc = (g.Dl + nchunks - 1) / nchunks;
    if (zc < 8) zc = 8;
    if (s->zc_override > 0) zc = s->zc_override;
    if (zc > g.Dl) zc = g.Dl;
    nchunks = (g.Dl + zc - 1) / zc;
    A.zc = zc;
    dim3 grid(A.ntx * nty, nchunks, nb);

    float* cur[3]; float* alt[3];
    for (int i = 0; i < nb; ++i) {
        cur[i] = h->field[fid[i]]; alt[i] = h->scratch[i];
        A.rhs[i] = h->field[pid[i]]; A.b[i] = b[i];
    }
    dim3 ge((2 * (g.W + 2) + 2 * g.H + 127) / 128, g.Dl + 2);
    for (int i = 0; i < nb; ++i) {                                        // edge/corner ghosts follow
        k_copy_edges<<<ge, 128, 0, h->stream>>>(g, cur[i], alt[i]);
        FL_CHECK_LAUNCH(h);
    }
    for (int k = 0; k < h->acc; ++k) {                                    // simulation.cpp:258
        for (int i = 0; i < nb; ++i) {
            if (int e = tl_get_map(h, cur[i], &A.map[i])) return e;
            A.dst[i] = alt[i];
        }
        if (solver == FLUID_SOLVER_RB) k_sweep_tiled<1><<<grid, TL_THREADS, smem, h->stream>>>(A);
        else                           k_sweep_tiled<0><<<grid, TL_THREADS, smem, h->stream>>>(A);
        FL_CHECK_LAUNCH(h);
        for (int i = 0; i < nb; ++i) { float* t = cur[i]; cur[i] = alt[i]; alt[i] = t; }
    }
    for (int i = 0; i < nb; ++i)
        if (cur[i] != h->field[fid[i]]) { h->scratch[i] = h->field[fid[i]]; h->field[fid[i]] = cur[i]; }
    *handled = true;
    return 0;
}
