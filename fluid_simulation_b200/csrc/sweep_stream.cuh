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
#include <cuda.h>
#include <map>
#include <cstdlib>

#define SW_BOXW   136                    // floats per box row: halo quad + 32 quads + halo quad
#define SW_AHEAD  3                      // TMA prefetch distance in planes
#define SW_NSLOT  (SW_AHEAD + 2)

struct alignas(64) StreamArgs {
    CUtensorMap map[3];                  // source field of each batched solve (box SW_BOXW x BR x 1)
    const float* rhs[3];
    float* dst[3];
    int b[3];
    const unsigned char* flags;
    Grid g;
    float a, c_recip;
    int zc;                              // planes per z-chunk
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
__device__ __forceinline__ void sw_fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ float4 sw_lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void   sw_sts4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

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

// edge bits of a thread's quad (which face ghosts its store also has to write)
#define SW_XLO 1u
#define SW_XHI 2u
#define SW_YLO 4u
#define SW_YHI 8u
#define SW_ZLO 16u
#define SW_ZHI 32u

// the setBounds epilogue for one finished quad (simulation.cpp:183-246), see fl_store_bounded:
// ghosts take the PRE-zero value, then solids (and, for velocity components, cells touching a
// solid) are zeroed.
__device__ __forceinline__ void sw_store(const Grid& g, float* __restrict__ out, long long c, const float4& raw,
                                         unsigned f4, int b, unsigned zmask4, unsigned edge)
{
    const unsigned t = f4 & zmask4;
    float4 v;
    v.x = (t & 0x000000ffu) ? 0.0f : raw.x;
    v.y = (t & 0x0000ff00u) ? 0.0f : raw.y;
    v.z = (t & 0x00ff0000u) ? 0.0f : raw.z;
    v.w = (t & 0xff000000u) ? 0.0f : raw.w;
    st4(out + c, v);
    if (edge) {
        if (edge & SW_XLO) out[c - 1] = (b == 1) ? -raw.x : raw.x;            // :189
        if (edge & SW_XHI) out[c + 4] = raw.w;                                // :191 outflow copy
        if (edge & (SW_YLO | SW_YHI)) {                                       // :198-201
            const float4 gy = (b == 2) ? make_float4(-raw.x, -raw.y, -raw.z, -raw.w) : raw;
            if (edge & SW_YLO) st4(out + c - g.pitch, gy);
            if (edge & SW_YHI) st4(out + c + g.pitch, gy);
        }
        if (edge & (SW_ZLO | SW_ZHI)) {                                       // :208-214
            const float4 gz = (b == 3) ? make_float4(-raw.x, -raw.y, -raw.z, -raw.w) : raw;
            if (edge & SW_ZLO) st4(out + c - g.plane, gz);
            if (edge & SW_ZHI) st4(out + c + g.plane, gz);
        }
    }
}

template <int TXW, int NP>
struct SwGeom {
    static constexpr int BR = 2 * NP + 2;                            // box rows: TY owned + 2 halo + 2 read-only
    static constexpr int TY = 2 * NP - 2;                            // owned rows per CTA
    static constexpr int BOXF = (SW_BOXW * BR + 31) / 32 * 32;       // floats per box, 128-byte multiple
    static constexpr int SLOTF = TXW * BOXF;                         // one plane of the CTA
    static constexpr int THREADS = 32 * TXW * NP;
    static constexpr size_t SMEM = (size_t)(SW_NSLOT + 2) * SLOTF * sizeof(float);
};

// MODE 0 = Jacobi, 1 = red-black (colour 0 = (x+y+z) even first)
template <int TXW, int NP, int MODE>
__global__ void __launch_bounds__(32 * TXW * NP, 1) k_sweep_stream(const __grid_constant__ StreamArgs A)
{
    using GE = SwGeom<TXW, NP>;
    constexpr int BOXF = GE::BOXF, SLOTF = GE::SLOTF;
    extern __shared__ __align__(128) float sw_mem[];
    __shared__ unsigned long long bars[SW_NSLOT];
    float* const uring = sw_mem;                                     // SW_NSLOT TMA slots (source field)
    float* const vbuf = sw_mem + SW_NSLOT * SLOTF;                   // 2 slots: plane after stage 0

    const Grid& g = A.g;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wx = warp % TXW, pr = warp / TXW;
    const int batch = blockIdx.z;
    const int y0 = 1 + blockIdx.x * GE::TY;                          // first owned row (odd)
    const int za = 1 + blockIdx.y * A.zc;
    const int zb = min(za + A.zc - 1, g.Dl);
    const int pa = za - 1, pb = zb + 1;                              // stage-0 plane range
    const int p_first = max(pa - 1, 0), p_last = min(pb + 1, g.Dl + 1);
    const float a = A.a, cr = A.c_recip;
    const int b = A.b[batch];
    const unsigned zmask4 = fl_zero_mask(b) * 0x01010101u;
    const float* __restrict__ rhs = A.rhs[batch];
    const unsigned char* __restrict__ flags = A.flags;
    float* __restrict__ out = A.dst[batch];

    auto slot_of = [&](int p) { return (p - p_first) % SW_NSLOT; };
    auto uslot = [&](int p) -> float* { return uring + slot_of(p) * SLOTF; };
    auto wait_plane = [&](int p) { sw_mbar_wait(&bars[slot_of(p)], ((p - p_first) / SW_NSLOT) & 1); };
    auto issue = [&](int p) {
        const int s = slot_of(p);
        sw_mbar_expect_tx(&bars[s], TXW * SW_BOXW * GE::BR * 4);
#pragma unroll
        for (int i = 0; i < TXW; ++i)
            sw_tma_load_3d(uring + s * SLOTF + i * BOXF, &A.map[batch], 128 * i, y0 - 2, p, &bars[s]);
    };

    if (tid == 0) {
        for (int i = 0; i < SW_NSLOT; ++i) sw_mbar_init(&bars[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0)
        for (int p = p_first; p <= min(pa + SW_AHEAD, p_last); ++p) issue(p);

    // this thread's quad: cells x0..x0+3 of rows yA (even) and yB = yA + 1 (odd)
    const int x0 = 1 + 4 * (32 * wx + lane);
    const int rA = 1 + 2 * pr;
    const int yA = y0 - 2 + rA, yB = yA + 1;
    const bool lane_in = x0 <= g.W;
    const bool intA = lane_in && yA >= 1 && yA <= g.H;               // row takes part in stage 0
    const bool intB = lane_in && yB >= 1 && yB <= g.H;
    const bool ownA = intA && pr >= 1;                               // row is stored by this CTA
    const bool ownB = intB && pr <= NP - 2;
    const int offA = wx * BOXF + rA * SW_BOXW + 4 * (lane + 1), offB = offA + SW_BOXW;
    const int offLA = wx * BOXF + rA * SW_BOXW + 3, offLB = offLA + SW_BOXW;      // cell left of lane 0
    const int offRA = offLA + 129, offRB = offLB + 129;                           // cell right of lane 31
    const unsigned exbits = (x0 == 1 ? SW_XLO : 0u) | (x0 + 3 == g.W ? SW_XHI : 0u);
    const unsigned eA = exbits | (yA == 1 ? SW_YLO : 0u) | (yA == g.H ? SW_YHI : 0u);
    const unsigned eB = exbits | (yB == 1 ? SW_YLO : 0u) | (yB == g.H ? SW_YHI : 0u);
    const long long idxA = fl_idx(g, x0, yA, 0), idxB = idxA + g.pitch;
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);

    // prologue: own column of planes pa-1 (behind) and pa (centre), rhs / flags of plane pa
    float4 z2A = zero4, z2B = zero4, mA = zero4, mB = zero4, cA, cB;
    if (pa - 1 >= 0) { wait_plane(pa - 1); mA = sw_lds4(uslot(pa - 1) + offA); mB = sw_lds4(uslot(pa - 1) + offB); }
    wait_plane(pa);
    cA = sw_lds4(uslot(pa) + offA); cB = sw_lds4(uslot(pa) + offB);
    float4 rcA = zero4, rcB = zero4, rpA = zero4, rpB = zero4;
    unsigned fcA = 0, fcB = 0, fpA = 0, fpB = 0;
    if (pa >= 1) {
        const long long o = (long long)pa * g.plane;
        if (intA) { rcA = ld4(rhs + idxA + o); fcA = *reinterpret_cast<const unsigned*>(flags + idxA + o); }
        if (intB) { rcB = ld4(rhs + idxB + o); fcB = *reinterpret_cast<const unsigned*>(flags + idxB + o); }
    }
    __syncthreads();        // slot of plane pa-1 is re-armed by the first iteration

    for (int p = pa; p <= pb; ++p) {
        if (tid == 0 && p + 1 + SW_AHEAD <= p_last) issue(p + 1 + SW_AHEAD);
        // rhs / flags of plane p+1 go out first: they are consumed one iteration later
        float4 rnA = zero4, rnB = zero4;
        unsigned fnA = 0, fnB = 0;
        if (p + 1 <= g.Dl) {
            const long long o = (long long)(p + 1) * g.plane;
            if (intA) { rnA = ld4(rhs + idxA + o); fnA = *reinterpret_cast<const unsigned*>(flags + idxA + o); }
            if (intB) { rnB = ld4(rhs + idxB + o); fnB = *reinterpret_cast<const unsigned*>(flags + idxB + o); }
        }
        float4 nA = zero4, nB = zero4;
        if (p + 1 <= p_last) {
            wait_plane(p + 1);
            const float* Un = uslot(p + 1);
            nA = sw_lds4(Un + offA); nB = sw_lds4(Un + offB);
        }
        const float* Up = uslot(p);
        const int E = (g.zlo + p) & 1;                               // (z_global(p) + 1) & 1

        if (MODE == 0) {
            // ---- Jacobi: relax all four lanes of plane p from the old planes p-1, p, p+1 ----------
            if (p >= za && p <= zb) {
                const float4 ymA = sw_lds4(Up + offA - SW_BOXW), ypB = sw_lds4(Up + offB + SW_BOXW);
                float LA = __shfl_up_sync(0xffffffffu, cA.w, 1), RA = __shfl_down_sync(0xffffffffu, cA.x, 1);
                float LB = __shfl_up_sync(0xffffffffu, cB.w, 1), RB = __shfl_down_sync(0xffffffffu, cB.x, 1);
                if (lane == 0) { LA = Up[offLA]; LB = Up[offLB]; }
                if (lane == 31) { RA = Up[offRA]; RB = Up[offRB]; }
                const unsigned zbits = (g.zlo + p - 1 == 1 ? SW_ZLO : 0u) | (g.zlo + p - 1 == g.D ? SW_ZHI : 0u);
                const long long o = (long long)p * g.plane;
                if (ownA) {
                    float4 e = cA, f = cA;
                    sw_relax2<0>(e, LA, RA, cB, ymA, nA, mA, rcA, a, cr);
                    sw_relax2<1>(f, LA, RA, cB, ymA, nA, mA, rcA, a, cr);
                    sw_store(g, out, idxA + o, make_float4(e.x, f.y, e.z, f.w), fcA, b, zmask4, eA | zbits);
                }
                if (ownB) {
                    float4 e = cB, f = cB;
                    sw_relax2<0>(e, LB, RB, ypB, cA, nB, mB, rcB, a, cr);
                    sw_relax2<1>(f, LB, RB, ypB, cA, nB, mB, rcB, a, cr);
                    sw_store(g, out, idxB + o, make_float4(e.x, f.y, e.z, f.w), fcB, b, zmask4, eB | zbits);
                }
            }
        } else {
            // ---- stage 0: colour-0 cells of plane p, in registers ----------------------------------
            if (p >= 1 && p <= g.Dl) {
                const float4 ymA = sw_lds4(Up + offA - SW_BOXW), ypB = sw_lds4(Up + offB + SW_BOXW);
                if (E == 0) {
                    float L = __shfl_up_sync(0xffffffffu, cA.w, 1), R = __shfl_down_sync(0xffffffffu, cB.x, 1);
                    if (lane == 0) L = Up[offLA];
                    if (lane == 31) R = Up[offRB];
                    if (intA) sw_relax2<0>(cA, L, 0.0f, cB, ymA, nA, mA, rcA, a, cr);
                    if (intB) sw_relax2<1>(cB, 0.0f, R, ypB, cA, nB, mB, rcB, a, cr);
                } else {
                    float R = __shfl_down_sync(0xffffffffu, cA.x, 1), L = __shfl_up_sync(0xffffffffu, cB.w, 1);
                    if (lane == 31) R = Up[offRA];
                    if (lane == 0) L = Up[offLB];
                    if (intA) sw_relax2<1>(cA, 0.0f, R, cB, ymA, nA, mA, rcA, a, cr);
                    if (intB) sw_relax2<0>(cB, L, 0.0f, ypB, cA, nB, mB, rcB, a, cr);
                }
            }
            // publish plane p (colour 0 new, colour 1 old) for the row / warp neighbours
            float* Vp = vbuf + (p & 1) * SLOTF;
            sw_sts4(Vp + offA, cA); sw_sts4(Vp + offB, cB);
            if (lane == 0) {
                if (wx > 0) { Vp[offRA - BOXF] = cA.x; Vp[offRB - BOXF] = cB.x; }
                else        { Vp[offLA] = Up[offLA];   Vp[offLB] = Up[offLB]; }      // x = 0 ghosts
            }
            if (lane == 31) {
                if (wx + 1 < TXW) { Vp[offLA + BOXF] = cA.w; Vp[offLB + BOXF] = cB.w; }
                else              { Vp[offRA] = Up[offRA];   Vp[offRB] = Up[offRB]; } // x = W+1 ghosts
            }
            // ---- stage 1: colour-1 cells of plane p-1 from colour-0 values, then setBounds ---------
            const int q = p - 1;
            if (q >= za && q <= zb) {
                const float* Vq = vbuf + (q & 1) * SLOTF;            // published (and synced) last iteration
                const float4 ymA = sw_lds4(Vq + offA - SW_BOXW), ypB = sw_lds4(Vq + offB + SW_BOXW);
                if (E == 0) {
                    float L = __shfl_up_sync(0xffffffffu, mA.w, 1), R = __shfl_down_sync(0xffffffffu, mB.x, 1);
                    if (lane == 0) L = Vq[offLA];
                    if (lane == 31) R = Vq[offRB];
                    if (ownA) sw_relax2<0>(mA, L, 0.0f, mB, ymA, cA, z2A, rpA, a, cr);
                    if (ownB) sw_relax2<1>(mB, 0.0f, R, ypB, mA, cB, z2B, rpB, a, cr);
                } else {
                    float R = __shfl_down_sync(0xffffffffu, mA.x, 1), L = __shfl_up_sync(0xffffffffu, mB.w, 1);
                    if (lane == 31) R = Vq[offRA];
                    if (lane == 0) L = Vq[offLB];
                    if (ownA) sw_relax2<1>(mA, 0.0f, R, mB, ymA, cA, z2A, rpA, a, cr);
                    if (ownB) sw_relax2<0>(mB, L, 0.0f, ypB, mA, cB, z2B, rpB, a, cr);
                }
                const int gzq = g.zlo + q - 1;
                const unsigned zbits = (gzq == 1 ? SW_ZLO : 0u) | (gzq == g.D ? SW_ZHI : 0u);
                const long long o = (long long)q * g.plane;
                if (ownA) sw_store(g, out, idxA + o, mA, fpA, b, zmask4, eA | zbits);
                if (ownB) sw_store(g, out, idxB + o, mB, fpB, b, zmask4, eB | zbits);
            }
        }
        // advance the register column by one plane
        z2A = mA; z2B = mB; mA = cA; mB = cB; cA = nA; cB = nB;
        rpA = rcA; rpB = rcB; rcA = rnA; rcB = rnB;
        fpA = fcA; fpB = fcB; fcA = fnA; fcB = fnB;
        sw_fence_proxy_async();        // generic-proxy reads of the slot the next TMA overwrites
        __syncthreads();
    }
}

// ---- host side -------------------------------------------------------------------------------
typedef CUresult (*sw_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                 CUtensorMapFloatOOBfill);

struct StreamState {
    sw_encode_fn encode = nullptr;
    std::map<std::pair<const float*, int>, CUtensorMap> maps;   // (field buffer, box rows) -> 3-D tensor map
    bool attr_set = false;
    int zc_override = 0;
    int np_override = 0;
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
        h->tiled = s;
    }
    return (StreamState*)h->tiled;
}

inline void fl_tiled_release(fluid_handle* h)
{
    delete (StreamState*)h->tiled;
    h->tiled = nullptr;
}

inline int sw_get_map(fluid_handle* h, const float* buf, int box_rows, CUtensorMap* out)
{
    StreamState* s = sw_state(h);
    auto key = std::make_pair(buf, box_rows);
    auto it = s->maps.find(key);
    if (it == s->maps.end()) {
        const Grid& g = h->g;
        CUtensorMap m;
        cuuint64_t dims[3] = { (cuuint64_t)g.pitch, (cuuint64_t)g.rows, (cuuint64_t)(g.Dl + 2) };
        cuuint64_t strides[2] = { (cuuint64_t)g.pitch * 4, (cuuint64_t)g.plane * 4 };
        cuuint32_t box[3] = { SW_BOXW, (cuuint32_t)box_rows, 1 };
        cuuint32_t estr[3] = { 1, 1, 1 };
        CUresult r = s->encode(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)buf, dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fl_fail(h, FLUID_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
        it = s->maps.emplace(key, m).first;
    }
    *out = it->second;
    return 0;
}

// z-chunking: CTAs are (row tiles) x (z chunks) x (batched fields), one CTA per SM.  Pick the chunk
// count whose last wave is best filled, discounted by the 4 extra planes each chunk re-reads.
inline int sw_pick_chunks(int ctas_per_chunk, int Dl, int sms)
{
    int best = 1; double best_eff = -1.0;
    for (int n = 1; n <= Dl; ++n) {
        const int zc = (Dl + n - 1) / n;
        if (n > 1 && zc < 6) break;
        const int nn = (Dl + zc - 1) / zc;
        const long long total = (long long)ctas_per_chunk * nn;
        const long long waves = (total + sms - 1) / sms;
        // time ~ waves * (zc + 4) planes; useful work = Dl planes per row tile
        const double eff = (double)Dl / ((double)waves * sms / ctas_per_chunk * (zc + 4));
        if (eff > best_eff + 1e-9) { best_eff = eff; best = nn; }
        if (total > 8LL * sms) break;
    }
    return best;
}

template <int TXW, int NP>
inline int sw_run(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid, float a, float c,
                  bool jacobi)
{
    using GE = SwGeom<TXW, NP>;
    const Grid& g = h->g;
    StreamState* s = sw_state(h);
    auto kj = k_sweep_stream<TXW, NP, 0>;
    auto kr = k_sweep_stream<TXW, NP, 1>;
    FL_CUDA(h, cudaFuncSetAttribute(kj, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GE::SMEM));
    FL_CUDA(h, cudaFuncSetAttribute(kr, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GE::SMEM));
    StreamArgs A;
    A.g = g; A.a = a; A.c_recip = 1.0f / c; A.flags = h->cflags;         // simulation.cpp:257
    const int nty = (g.H + GE::TY - 1) / GE::TY;
    int nchunks = sw_pick_chunks(nty * nb, g.Dl, h->sm_count);
    int zc = (g.Dl + nchunks - 1) / nchunks;
    if (s->zc_override > 0) zc = s->zc_override;
    if (zc > g.Dl) zc = g.Dl;
    nchunks = (g.Dl + zc - 1) / zc;
    A.zc = zc;
    dim3 grid(nty, nchunks, nb);

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
            if (int e = sw_get_map(h, cur[i], GE::BR, &A.map[i])) return e;
            A.dst[i] = alt[i];
        }
        if (jacobi) kj<<<grid, GE::THREADS, GE::SMEM, h->stream>>>(A);
        else        kr<<<grid, GE::THREADS, GE::SMEM, h->stream>>>(A);
        FL_CHECK_LAUNCH(h);
        for (int i = 0; i < nb; ++i) { float* t = cur[i]; cur[i] = alt[i]; alt[i] = t; }
    }
    for (int i = 0; i < nb; ++i)
        if (cur[i] != h->field[fid[i]]) { h->scratch[i] = h->field[fid[i]]; h->field[fid[i]] = cur[i]; }
    return 0;
}

// linearSolver through the streaming kernel; *handled = false leaves the solve to the simple kernels
inline int fl_tiled_solve(fluid_handle* h, int nb, const int* b, const int* fid, const int* pid,
                          float a, float c, bool* handled)
{
    *handled = false;
    const Grid& g = h->g;
    const unsigned solver = h->flags & FLUID_SOLVER_MASK;
    if (solver != FLUID_SOLVER_RB && solver != FLUID_SOLVER_JACOBI) return 0;
    if (h->nranks > 1) return 0;                         // slab mode exchanges halos per half sweep (simple path)
    if ((g.W & 3) || g.W > 512) return 0;                // quads must not straddle the x = W+1 ghost
    if (!(h->flags & FLUID_TILED_ALWAYS) && (g.W < 64 || g.H < 8 || g.Dl < 4)) return 0;
    StreamState* s = sw_state(h);
    if (!s->encode) return fl_fail(h, FLUID_E_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
    const bool jac = solver == FLUID_SOLVER_JACOBI;
    int e;
    if (g.W <= 128)      e = sw_run<1, 16>(h, nb, b, fid, pid, a, c, jac);
    else if (g.W <= 256) e = sw_run<2, 9>(h, nb, b, fid, pid, a, c, jac);
    else                 e = sw_run<4, 4>(h, nb, b, fid, pid, a, c, jac);
    if (e) return e;
    *handled = true;
    return 0;
}
