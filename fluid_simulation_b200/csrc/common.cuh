// common.cuh -- grid geometry, HBM layout and small device helpers shared by all kernels.
//
// HBM layout (DESIGN.md "Data layout"): every field is a pitched SoA float volume
//   index(x, y, zl) = XOFF + x + y*pitch + zl*plane,   x in 0..W+1, y in 0..H+1, zl in 0..Dl+1
// with XOFF = 3 so that the first interior cell x=1 sits on a 16-byte boundary (column 4) and
// pitch = round_up(W + 2 + XOFF, 4): interior x-rows are float4-aligned and every row stride is a
// multiple of 16 B (the TMA requirement).  The reference layout (x + y*(W+2) + z*(W+2)*(H+2),
// simulation.h:9) only exists in the pack/unpack kernels at the host boundary.
// zl is the LOCAL plane index of a z-slab: global z = zlo + zl - 1 (zlo = 1, Dl = D on one GPU).
// Every volume is allocated with planes zl = -FL_ZPAD .. Dl+1+FL_ZPAD (the field pointer addresses plane 0).
// A slab keeps FL_HALO = 4 halo planes per interior side: one fused red-black sweep needs TWO (the colour-0
// update of the neighbour's boundary plane is recomputed locally from the second halo plane), the pass
// that fuses two sweeps needs four.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define FL_XOFF 3
#define FL_HALO 4    // halo planes a z-slab keeps per interior side: one temporally blocked pass (two red-black
                     // sweeps, sweep_pair_kernel.cuh) reads its input four planes beyond the planes it owns
#define FL_ZPAD 4    // planes allocated below local plane 0 and above plane Dl+1 (local planes -4 .. Dl+5)

// per-cell flag byte, built once from obs (k_build_flags)
#define FL_SOLID   0x01u   // obs == 1.0f                       (simulation.cpp:222, 301, 326, 375)
#define FL_NOSLIP  0x02u   // fluid cell with an in-domain solid 6-neighbour (simulation.cpp:232-238)
#define FL_OPEN_XP 0x04u   // i+1 <= W  && obs(i+1) == 0.0f     (simulation.cpp:307, 329)
#define FL_OPEN_XM 0x08u   // i-1 >= 1  && obs(i-1) == 0.0f
#define FL_OPEN_YP 0x10u
#define FL_OPEN_YM 0x20u
#define FL_OPEN_ZP 0x40u
#define FL_OPEN_ZM 0x80u

struct Grid {
    int W, H, D;          // global interior size
    int Dl, zlo;          // local slab: interior planes and global z of local plane 1
    int pitch;            // floats per x-row
    int rows;             // H + 2
    long long plane;      // pitch * rows
    long long cells;      // plane * (Dl + 2 + 2 * FL_ZPAD)   (allocated floats per field)
    int zl_lo, zl_hi;     // local planes that hold real data: 0 / Dl+1 at the tunnel ends, 1-FL_HALO / Dl+FL_HALO inside
    int nxq;              // ceil(W / 4) x-quads per row
};

__host__ __device__ __forceinline__ long long fl_idx(const Grid& g, int x, int y, int zl)
{
    return (long long)FL_XOFF + x + (long long)y * g.pitch + (long long)zl * g.plane;
}

// the 7-point update of linearSolver (simulation.cpp:263-269); neighbour sum strictly left to
// right: x+1, x-1, y+1, y-1, z+1, z-1
__device__ __forceinline__ float fl_relax(float rhs, float a, float c_recip, float xp, float xm,
                                          float yp, float ym, float zp, float zm)
{
    return (rhs + a * (((((xp + xm) + yp) + ym) + zp) + zm)) * c_recip;
}

// velocity component b in {1,2,3} masks solids and no-slip cells, scalars (b == 0) only solids
__device__ __forceinline__ unsigned fl_zero_mask(int b) { return b == 0 ? FL_SOLID : (FL_SOLID | FL_NOSLIP); }

// thread -> (x-quad, y, zl) mapping used by all "one thread = 4 consecutive x cells" kernels:
// blockDim = (bx, by), gridDim = (ceil(nxq/bx), ceil(H/by), Dl [* batch])
struct QuadPos { int x0, y, zl; bool valid; };

__device__ __forceinline__ QuadPos fl_quad_pos(const Grid& g, int zl)
{
    QuadPos p;
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    p.x0 = 1 + 4 * q;
    p.y  = 1 + blockIdx.y * blockDim.y + threadIdx.y;
    p.zl = zl;
    p.valid = (q < g.nxq) && (p.y <= g.H);
    return p;
}

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void   st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ uchar4 ldf4(const unsigned char* p) { return *reinterpret_cast<const uchar4*>(p); }

// The setBounds epilogue fused into every kernel that writes a field (simulation.cpp:183-246).
// raw[i] is the value the reference has in the interior cell BEFORE setBounds runs; the face
// ghosts copy/negate that PRE-zero value (:187-215), then the cell itself is zeroed if it is
// solid (:219-223) or, for velocity components, touches a solid (:227-245).  Edge/corner ghosts
// are never written.  In a z-slab the z ghosts exist only on the slabs that own z=1 / z=D.
__device__ __forceinline__ void fl_store_bounded(const Grid& g, float* __restrict__ out,
                                                 const unsigned char* __restrict__ flags,
                                                 int b, int x0, int y, int zl, const float raw[4])
{
    const long long c = fl_idx(g, x0, y, zl);
    const uchar4 f4 = ldf4(flags + c);
    const unsigned fl[4] = { f4.x, f4.y, f4.z, f4.w };
    const unsigned zm = fl_zero_mask(b);
    float v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = (fl[i] & zm) ? 0.0f : raw[i];

    const int nvalid = min(4, g.W - x0 + 1);
    if (nvalid == 4) {
        st4(out + c, make_float4(v[0], v[1], v[2], v[3]));
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) if (i < nvalid) out[c + i] = v[i];
    }
    // x faces (:187-192): x=0 negates only for b==1, x=W+1 is a plain copy (outflow)
    if (x0 == 1) out[c - 1] = (b == 1) ? -raw[0] : raw[0];
#pragma unroll
    for (int i = 0; i < 4; ++i) if (x0 + i == g.W) out[c + i + 1] = raw[i];
    // y faces (:196-202)
    if (y == 1 || y == g.H) {
        float gy[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) gy[i] = (b == 2) ? -raw[i] : raw[i];
        if (y == 1) {
            float* o = out + c - g.pitch;
            if (nvalid == 4) st4(o, make_float4(gy[0], gy[1], gy[2], gy[3]));
            else { for (int i = 0; i < nvalid; ++i) o[i] = gy[i]; }
        }
        if (y == g.H) {
            float* o = out + c + g.pitch;
            if (nvalid == 4) st4(o, make_float4(gy[0], gy[1], gy[2], gy[3]));
            else { for (int i = 0; i < nvalid; ++i) o[i] = gy[i]; }
        }
    }
    // z faces (:206-215), only on the slab that owns the global boundary plane
    const int gz = g.zlo + zl - 1;
    if (gz == 1 || gz == g.D) {
        float gzv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) gzv[i] = (b == 3) ? -raw[i] : raw[i];
        if (gz == 1) {
            float* o = out + c - g.plane;
            if (nvalid == 4) st4(o, make_float4(gzv[0], gzv[1], gzv[2], gzv[3]));
            else { for (int i = 0; i < nvalid; ++i) o[i] = gzv[i]; }
        }
        if (gz == g.D) {
            float* o = out + c + g.plane;
            if (nvalid == 4) st4(o, make_float4(gzv[0], gzv[1], gzv[2], gzv[3]));
            else { for (int i = 0; i < nvalid; ++i) o[i] = gzv[i]; }
        }
    }
}

// ---- lean form of the same epilogue for FULL quads (x0 + 3 <= W) ------------------------------------
// The face tests are folded into edge bits that a kernel writing several fields computes once per
// quad; away from the faces (edge == 0) the epilogue is four selects and one 16-byte store.
#define FL_XLO 1u
#define FL_XHI 2u
#define FL_YLO 4u
#define FL_YHI 8u
#define FL_ZLO 16u
#define FL_ZHI 32u

__device__ __forceinline__ unsigned fl_edge_bits(const Grid& g, int x0, int y, int zl)
{
    const int gz = g.zlo + zl - 1;
    return (x0 == 1 ? FL_XLO : 0u) | (x0 + 3 == g.W ? FL_XHI : 0u) | (y == 1 ? FL_YLO : 0u) |
           (y == g.H ? FL_YHI : 0u) | (gz == 1 ? FL_ZLO : 0u) | (gz == g.D ? FL_ZHI : 0u);
}

// f4 = the four flag bytes of the quad (little endian: lane 0 in the low byte)
__device__ __forceinline__ void fl_store_quad(const Grid& g, float* __restrict__ out, long long c, const float4& raw,
                                              unsigned f4, int b, unsigned edge)
{
    const unsigned t = f4 & (fl_zero_mask(b) * 0x01010101u);
    float4 v;
    v.x = (t & 0x000000ffu) ? 0.0f : raw.x;
    v.y = (t & 0x0000ff00u) ? 0.0f : raw.y;
    v.z = (t & 0x00ff0000u) ? 0.0f : raw.z;
    v.w = (t & 0xff000000u) ? 0.0f : raw.w;
    st4(out + c, v);
    if (edge) {
        if (edge & FL_XLO) out[c - 1] = (b == 1) ? -raw.x : raw.x;            // simulation.cpp:189
        if (edge & FL_XHI) out[c + 4] = raw.w;                                // :191 outflow copy
        if (edge & (FL_YLO | FL_YHI)) {                                       // :198-201
            const float4 gy = (b == 2) ? make_float4(-raw.x, -raw.y, -raw.z, -raw.w) : raw;
            if (edge & FL_YLO) st4(out + c - g.pitch, gy);
            if (edge & FL_YHI) st4(out + c + g.pitch, gy);
        }
        if (edge & (FL_ZLO | FL_ZHI)) {                                       // :208-214
            const float4 gz = (b == 3) ? make_float4(-raw.x, -raw.y, -raw.z, -raw.w) : raw;
            if (edge & FL_ZLO) st4(out + c - g.plane, gz);
            if (edge & FL_ZHI) st4(out + c + g.plane, gz);
        }
    }
}
