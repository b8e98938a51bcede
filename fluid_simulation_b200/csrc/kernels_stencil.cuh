// kernels_stencil.cuh -- simple global-memory versions of the stencil kernels (one thread = one
// x-quad, float4 rows, L1/L2 provide the neighbour reuse).  They are the reference-exact
// baseline that the TMA/shared-memory tiled sweeps (kernels_tiled.cuh) are A/B-checked against,
// and they handle every grid shape.  All arithmetic is compiled with --fmad=false and keeps the
// reference's operand order so results are bit-identical to the CPU code.
#pragma once
#include "common.cuh"
#include "recip_div.h"

struct Quad7 { float c[4], xm, xp, yp[4], ym[4], zp[4], zm[4]; };

__device__ __forceinline__ void fl_load_quad7(const Grid& g, const float* __restrict__ f,
                                              long long c, Quad7& q)
{
    float4 v = ld4(f + c);            q.c[0] = v.x;  q.c[1] = v.y;  q.c[2] = v.z;  q.c[3] = v.w;
    q.xm = f[c - 1];                  q.xp = f[c + 4];
    v = ld4(f + c + g.pitch);         q.yp[0] = v.x; q.yp[1] = v.y; q.yp[2] = v.z; q.yp[3] = v.w;
    v = ld4(f + c - g.pitch);         q.ym[0] = v.x; q.ym[1] = v.y; q.ym[2] = v.z; q.ym[3] = v.w;
    v = ld4(f + c + g.plane);         q.zp[0] = v.x; q.zp[1] = v.y; q.zp[2] = v.z; q.zp[3] = v.w;
    v = ld4(f + c - g.plane);         q.zm[0] = v.x; q.zm[1] = v.y; q.zm[2] = v.z; q.zm[3] = v.w;
}

// ---- Jacobi sweep + fused setBounds: dst = bounds(relax(src)), simulation.cpp:259-271 --------
// Batched over `nb` fields (blockIdx.z = batch * Dl + plane) so the three velocity diffusions,
// which share a and c (simulation.cpp:282-283), go out as one launch.
struct SweepBatch { const float* src[3]; const float* rhs[3]; float* dst[3]; int b[3]; };

__global__ void __launch_bounds__(256) k_sweep_jacobi(Grid g, SweepBatch sb,
                                                      const unsigned char* __restrict__ flags,
                                                      float a, float c_recip)
{
    const int batch = blockIdx.z / g.Dl;
    QuadPos p = fl_quad_pos(g, 1 + blockIdx.z % g.Dl);
    if (!p.valid) return;
    const float* __restrict__ src = sb.src[batch];
    const float* __restrict__ rhs = sb.rhs[batch];
    const long long c = fl_idx(g, p.x0, p.y, p.zl);
    Quad7 q; fl_load_quad7(g, src, c, q);
    float4 r4 = ld4(rhs + c);
    const float r[4] = { r4.x, r4.y, r4.z, r4.w };
    float raw[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float xp = (i == 3) ? q.xp : q.c[i + 1];
        float xm = (i == 0) ? q.xm : q.c[i - 1];
        raw[i] = fl_relax(r[i], a, c_recip, xp, xm, q.yp[i], q.ym[i], q.zp[i], q.zm[i]);
    }
    fl_store_bounded(g, sb.dst[batch], flags, sb.b[batch], p.x0, p.y, p.zl, raw);
}

// ---- red-black half sweep, in place, no bounds (they run after the black half) ---------------
// colour 0 = cells with (i+j+l) even, global coordinates.  Only cells of `colour` change; the
// other two lanes of the quad are written back unchanged (nobody else writes them in this pass).
// Planes zl_first .. zl_first+nplanes-1 are swept: 1..Dl normally; a z-slab runs colour 0 also on its
// first halo plane (0 / Dl+1 where that plane is interior to the tunnel) so that the colour-1 pass sees
// the neighbour's colour-0 update without an exchange in the middle of the sweep.
__global__ void __launch_bounds__(256) k_sweep_rb_half(Grid g, SweepBatch sb, float a, float c_recip,
                                                       int colour, int zl_first, int nplanes)
{
    const int batch = blockIdx.z / nplanes;
    QuadPos p = fl_quad_pos(g, zl_first + blockIdx.z % nplanes);
    if (!p.valid) return;
    float* __restrict__ f = sb.dst[batch];
    const float* __restrict__ rhs = sb.rhs[batch];
    const long long c = fl_idx(g, p.x0, p.y, p.zl);
    Quad7 q; fl_load_quad7(g, f, c, q);
    float4 r4 = ld4(rhs + c);
    const float r[4] = { r4.x, r4.y, r4.z, r4.w };
    const int gz = g.zlo + p.zl - 1;
    const int par0 = (p.x0 + p.y + gz) & 1;       // parity of lane 0
    float out[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float xp = (i == 3) ? q.xp : q.c[i + 1];
        float xm = (i == 0) ? q.xm : q.c[i - 1];
        float v = fl_relax(r[i], a, c_recip, xp, xm, q.yp[i], q.ym[i], q.zp[i], q.zm[i]);
        out[i] = (((par0 + i) & 1) == colour) ? v : q.c[i];
    }
    const int nvalid = min(4, g.W - p.x0 + 1);
    if (nvalid == 4) st4(f + c, make_float4(out[0], out[1], out[2], out[3]));
    else for (int i = 0; i < nvalid; ++i) f[c + i] = out[i];
}

// ---- project part 1: divergence, p = 0, setBounds(0,div), setBounds(0,p) (:295-319) ----------
__global__ void __launch_bounds__(256) k_divergence(Grid g, const float* __restrict__ vx,
                                                    const float* __restrict__ vy,
                                                    const float* __restrict__ vz,
                                                    const unsigned char* __restrict__ flags,
                                                    float* __restrict__ div, float* __restrict__ p,
                                                    float h)
{
    QuadPos q = fl_quad_pos(g, 1 + blockIdx.z);
    if (!q.valid) return;
    const long long c = fl_idx(g, q.x0, q.y, q.zl);
    const uchar4 f4 = ldf4(flags + c);
    const unsigned fl[4] = { f4.x, f4.y, f4.z, f4.w };
    const float4 xc = ld4(vx + c);
    const float vxq[6] = { vx[c - 1], xc.x, xc.y, xc.z, xc.w, vx[c + 4] };
    const float4 a = ld4(vy + c + g.pitch), bq = ld4(vy + c - g.pitch);
    const float4 cz = ld4(vz + c + g.plane), dz = ld4(vz + c - g.plane);
    const float yp[4] = { a.x, a.y, a.z, a.w }, ym[4] = { bq.x, bq.y, bq.z, bq.w };
    const float zp[4] = { cz.x, cz.y, cz.z, cz.w }, zm[4] = { dz.x, dz.y, dz.z, dz.w };
    float raw[4], zero[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float s = 0.0f;                                           // :306-312, same order
        if (fl[i] & FL_OPEN_XP) s += vxq[i + 2];
        if (fl[i] & FL_OPEN_XM) s -= vxq[i];
        if (fl[i] & FL_OPEN_YP) s += yp[i];
        if (fl[i] & FL_OPEN_YM) s -= ym[i];
        if (fl[i] & FL_OPEN_ZP) s += zp[i];
        if (fl[i] & FL_OPEN_ZM) s -= zm[i];
        raw[i] = (fl[i] & FL_SOLID) ? 0.0f : -0.5f * h * s;       // :301-305, :314
    }
    fl_store_bounded(g, div, flags, 0, q.x0, q.y, q.zl, raw);
    fl_store_bounded(g, p, flags, 0, q.x0, q.y, q.zl, zero);
}

// one-sided / central pressure difference (simulation.cpp:328-336 and the y, z copies) in select
// form: both open -> (p+ - p-)/(2h); only + -> (p+ - p)/h; only - -> (p - p-)/h; none -> 0.
// One division per axis instead of three divergent ones (the kernel was issue bound on them).
// POW2: h is a power of two (every cubic 2^k grid: h = 1/cbrtf(W*H*D) = 2^-k), so x / h == x * (1/h) and
// x / (2h) == x * (1/(2h)) bit for bit for EVERY float x (exact scaling; zeros keep their sign, denormals,
// infinities and NaNs behave the same): s1 = 1/h, s2 = 1/(2h) and the quotient is one FMUL.  The IEEE
// division it replaces is ~15 instructions on its fast path and a ~100-instruction subroutine whenever the
// numerator is zero or denormal -- which the far field of the pressure solve is full of.  Otherwise
// s1 = h, s2 = 2h and the division stays.
// Any other h: FL_GRAD_RECIP divides by multiplying with the correctly rounded reciprocal and one FMA
// correction (recip_div.h; verified bit-identical to the division for the launch's two divisors on the host
// before the first launch); FL_GRAD_DIV keeps the division (A/B, or a divisor that failed the verification).
enum { FL_GRAD_DIV = 0, FL_GRAD_POW2 = 1, FL_GRAD_RECIP = 2 };
struct GradScale { float s1, s2, r1, r2; };          // h, 2h and (POW2, RECIP) their reciprocals

template <int MODE>
__device__ __forceinline__ float fl_axis_grad(bool po, bool mo, float pp, float pc, float pm, const GradScale& k)
{
    const float hi = po ? pp : pc, lo = mo ? pm : pc;
    const bool both = po && mo;
    const float n = hi - lo;
    float q;
    if (MODE == FL_GRAD_POW2)       q = n * (both ? k.r2 : k.r1);
    else if (MODE == FL_GRAD_RECIP) q = fl_recip_div(n, both ? k.s2 : k.s1, both ? k.r2 : k.r1);
    else                            q = n / (both ? k.s2 : k.s1);
    return (po || mo) ? q : 0.0f;
}

// flag bit `m` of lane i in the packed flag word of a quad (one LOP3 with a predicate result)
#define FL_QBIT(f4, i, m) (((f4) & ((unsigned)(m) << (8 * (i)))) != 0u)

// ---- project part 3: v -= grad p on fluid cells, then setBounds(1..3) (:322-361) -------------
template <int MODE>
__global__ void __launch_bounds__(256) k_gradient(Grid g, const float* __restrict__ p,
                                                  const unsigned char* __restrict__ flags,
                                                  float* __restrict__ vx, float* __restrict__ vy,
                                                  float* __restrict__ vz, GradScale sc)
{
    QuadPos qp = fl_quad_pos(g, 1 + blockIdx.z);
    if (!qp.valid) return;
    const long long c = fl_idx(g, qp.x0, qp.y, qp.zl);
    const unsigned f4 = *reinterpret_cast<const unsigned*>(flags + c);
    Quad7 q; fl_load_quad7(g, p, c, q);
    const float4 ax = ld4(vx + c), ay = ld4(vy + c), az = ld4(vz + c);
    float rx[4] = { ax.x, ax.y, ax.z, ax.w }, ry[4] = { ay.x, ay.y, ay.z, ay.w },
          rz[4] = { az.x, az.y, az.z, az.w };
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float pxp = (i == 3) ? q.xp : q.c[i + 1];
        const float pxm = (i == 0) ? q.xm : q.c[i - 1];
        const float gx = fl_axis_grad<MODE>(FL_QBIT(f4, i, FL_OPEN_XP), FL_QBIT(f4, i, FL_OPEN_XM), pxp, q.c[i], pxm, sc);
        const float gy = fl_axis_grad<MODE>(FL_QBIT(f4, i, FL_OPEN_YP), FL_QBIT(f4, i, FL_OPEN_YM), q.yp[i], q.c[i], q.ym[i], sc);
        const float gz = fl_axis_grad<MODE>(FL_QBIT(f4, i, FL_OPEN_ZP), FL_QBIT(f4, i, FL_OPEN_ZM), q.zp[i], q.c[i], q.zm[i], sc);
        const bool solid = FL_QBIT(f4, i, FL_SOLID);              // :326 solids keep their value
        rx[i] = solid ? rx[i] : rx[i] - gx;
        ry[i] = solid ? ry[i] : ry[i] - gy;
        rz[i] = solid ? rz[i] : rz[i] - gz;
    }
    if (qp.x0 + 3 <= g.W) {
        const unsigned edge = fl_edge_bits(g, qp.x0, qp.y, qp.zl);
        fl_store_quad(g, vx, c, make_float4(rx[0], rx[1], rx[2], rx[3]), f4, 1, edge);
        fl_store_quad(g, vy, c, make_float4(ry[0], ry[1], ry[2], ry[3]), f4, 2, edge);
        fl_store_quad(g, vz, c, make_float4(rz[0], rz[1], rz[2], rz[3]), f4, 3, edge);
    } else {                                                      // ragged last quad of a row
        fl_store_bounded(g, vx, flags, 1, qp.x0, qp.y, qp.zl, rx);
        fl_store_bounded(g, vy, flags, 2, qp.x0, qp.y, qp.zl, ry);
        fl_store_bounded(g, vz, flags, 3, qp.x0, qp.y, qp.zl, rz);
    }
}

// ---- advect(b, field, prev) + setBounds (simulation.cpp:367-424) -----------------------------
// `src` is the gather source (prev_field) as an array of per-slab base pointers: src_slab[r]
// holds global planes src_zlo[r]-1 .. ; on one GPU there is a single entry.  The own-cell
// velocity of component b comes from prev (:380-382), the other two from the current fields.
struct GatherSrc { const float* base[8]; int zlo[8]; int dl[8]; int n; };

// pointer to cell (0, 0) of global plane gz: the plane lives on the slab with zlo <= gz < zlo+dl
// (ghost planes 0 / D+1 on the end slabs)
__device__ __forceinline__ const float* fl_gather_plane(const Grid& g, const GatherSrc& s, int gz)
{
    int r = 0;
#pragma unroll 1
    for (int k = 0; k < s.n - 1; ++k) if (gz >= s.zlo[k] + s.dl[k]) r = k + 1;
    return s.base[r] + (FL_XOFF + (long long)(gz - s.zlo[r] + 1) * g.plane);
}

__device__ __forceinline__ float fl_clamp(float v, float lo, float hi)       // std::clamp
{
    return (v < lo) ? lo : (hi < v) ? hi : v;
}

__global__ void __launch_bounds__(256) k_advect(Grid g, int b, float* __restrict__ field,
                                                const float* __restrict__ prev_local, GatherSrc src,
                                                const float* __restrict__ vxf,
                                                const float* __restrict__ vyf,
                                                const float* __restrict__ vzf,
                                                const unsigned char* __restrict__ flags, float dt)
{
    QuadPos qp = fl_quad_pos(g, 1 + blockIdx.z);
    if (!qp.valid) return;
    const long long c = fl_idx(g, qp.x0, qp.y, qp.zl);
    const uchar4 f4 = ldf4(flags + c);
    const unsigned fl[4] = { f4.x, f4.y, f4.z, f4.w };
    const float4 ax = ld4((b == 1 ? prev_local : vxf) + c);
    const float4 ay = ld4((b == 2 ? prev_local : vyf) + c);
    const float4 az = ld4((b == 3 ? prev_local : vzf) + c);
    const float vx[4] = { ax.x, ax.y, ax.z, ax.w }, vy[4] = { ay.x, ay.y, ay.z, ay.w },
                vz[4] = { az.x, az.y, az.z, az.w };
    const int gz = g.zlo + qp.zl - 1;
    const float fW = (float)g.W, fH = (float)g.H, fD = (float)g.D;
    float raw[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        raw[i] = 0.0f;
        if ((fl[i] & FL_SOLID) || qp.x0 + i > g.W) continue;                 // :375-378
        float xb = (float)(qp.x0 + i) - dt * fW * vx[i];                     // :384-386
        float yb = (float)qp.y - dt * fH * vy[i];
        float zb = (float)gz - dt * fD * vz[i];
        xb = fl_clamp(xb, 0.5f, fW + 0.5f);                                  // :388-390
        yb = fl_clamp(yb, 0.5f, fH + 0.5f);
        zb = fl_clamp(zb, 0.5f, fD + 0.5f);
        const int i0 = (int)floorf(xb), j0 = (int)floorf(yb), k0 = (int)floorf(zb);   // :392-397
        const float sx = xb - (float)i0, sy = yb - (float)j0, sz = zb - (float)k0;    // :399-401
        // the 8 corners: two planes (k0, k0+1, possibly on different slabs), 2 x 2 cells in each
        const float* __restrict__ q0 = fl_gather_plane(g, src, k0) + (i0 + j0 * g.pitch);
        const float* __restrict__ q1 = fl_gather_plane(g, src, k0 + 1) + (i0 + j0 * g.pitch);
        const float c000 = __ldg(q0), c100 = __ldg(q0 + 1), c010 = __ldg(q0 + g.pitch), c110 = __ldg(q0 + g.pitch + 1);
        const float c001 = __ldg(q1), c101 = __ldg(q1 + 1), c011 = __ldg(q1 + g.pitch), c111 = __ldg(q1 + g.pitch + 1);
        const float c00 = c000 * (1.0f - sx) + c100 * sx;                    // :412-415
        const float c01 = c001 * (1.0f - sx) + c101 * sx;
        const float c10 = c010 * (1.0f - sx) + c110 * sx;
        const float c11 = c011 * (1.0f - sx) + c111 * sx;
        const float c0 = c00 * (1.0f - sy) + c10 * sy;                       // :417-418
        const float c1 = c01 * (1.0f - sy) + c11 * sy;
        raw[i] = c0 * (1.0f - sz) + c1 * sz;                                 // :420
    }
    if (qp.x0 + 3 <= g.W)
        fl_store_quad(g, field, c, make_float4(raw[0], raw[1], raw[2], raw[3]),
                      *reinterpret_cast<const unsigned*>(flags + c), b, fl_edge_bits(g, qp.x0, qp.y, qp.zl));
    else
        fl_store_bounded(g, field, flags, b, qp.x0, qp.y, qp.zl, raw);
}

// ld.global.nc as a volatile asm: the compiler keeps volatile asm statements in program order, so a batch of
// these loads is issued as written and cannot be sunk behind a later store (it does that with __ldg).
__device__ __forceinline__ float fl_ldnc(const float* p)
{
    float v;
    asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

// ---- advect, lanes along x, NY rows per thread -------------------------------------------------------
// The quad form above is bound by L1 and by issue (ncu: l1tex 89 %, issue 82 %): a warp's gather of
// "corner k of lane cell i" strides 16 B from lane to lane, so each of the 32 scalar loads of a thread
// touches four 128-byte lines, and a third of its instructions are 64-bit address arithmetic.  Here
// consecutive lanes sit on consecutive x cells: the eight gathers are contiguous across the warp wherever the
// back-trace is smooth (one or two lines per load), the own-cell loads and the store are plain coalesced
// scalars, every index is a 32-bit element offset (one IMAD.WIDE per address; the host uses this kernel only
// when the pitched volume has fewer than 2^31 cells), and a thread walks NY rows to amortise its set-up.
// SLAB = false: one GPU, every plane lives in src.base[0].
// Same statements, same operand order as k_advect: bit-identical results.
template <bool SLAB, int NY>
__global__ void __launch_bounds__(256) k_advect_rows(Grid g, int b, float* __restrict__ field,
                                                     const float* __restrict__ prev_local, GatherSrc src,
                                                     const float* __restrict__ vxf, const float* __restrict__ vyf,
                                                     const float* __restrict__ vzf,
                                                     const unsigned char* __restrict__ flags, float dt)
{
    const int x = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    const int ya = 1 + (blockIdx.y * blockDim.y + threadIdx.y) * NY;
    const int zl = 1 + blockIdx.z;
    if (x > g.W || ya > g.H) return;
    const unsigned pitch = (unsigned)g.pitch, plane = (unsigned)g.plane;
    const int gz = g.zlo + zl - 1;
    const float fW = (float)g.W, fH = (float)g.H, fD = (float)g.D;
    const float kx = dt * fW, ky = dt * fH, kz = dt * fD;                    // dt * width etc. (:384-386, left to right)
    const float hx = fW + 0.5f, hy = fH + 0.5f, hz = fD + 0.5f;
    const float* __restrict__ ux = b == 1 ? prev_local : vxf;                // :380-382
    const float* __restrict__ uy = b == 2 ? prev_local : vyf;
    const float* __restrict__ uz = b == 3 ? prev_local : vzf;
    const float* __restrict__ gsrc = src.base[0];
    const int zoff = 1 - src.zlo[0];
    const unsigned zm = fl_zero_mask(b);
    const bool xz_edge = x == 1 || x == g.W || gz == 1 || gz == g.D;
    const float fx = (float)x, fz = (float)gz;
    const unsigned c0 = FL_XOFF + (unsigned)x + (unsigned)ya * pitch + (unsigned)zl * plane;
    // Three phases without a branch between them, so that all NY flag / velocity loads and then all 8 * NY
    // gathers of a thread are in flight together (the row-by-row form was a chain of 3 * NY memory round trips).
    // Rows past H repeat row ya and are not stored; solid cells back-trace with zero velocity (their result is
    // discarded, :375-378) so that whatever their stored velocity is, the gather address stays inside the volume.
    unsigned cc[NY], fl[NY];
    float vx[NY], vy[NY], vz[NY];
#pragma unroll
    for (int j = 0; j < NY; ++j) {
        cc[j] = (ya + j <= g.H) ? c0 + (unsigned)j * pitch : c0;
        fl[j] = flags[cc[j]];
        vx[j] = ux[cc[j]]; vy[j] = uy[cc[j]]; vz[j] = uz[cc[j]];
    }
    float sx[NY], sy[NY], sz[NY], q[NY][8];
#pragma unroll
    for (int j = 0; j < NY; ++j) {
        const bool solid = fl[j] & FL_SOLID;
        const float fy = (float)((ya + j <= g.H) ? ya + j : ya);
        float xb = fx - kx * (solid ? 0.0f : vx[j]);                         // :384-386
        float yb = fy - ky * (solid ? 0.0f : vy[j]);
        float zb = fz - kz * (solid ? 0.0f : vz[j]);
        xb = fl_clamp(xb, 0.5f, hx);                                         // :388-390
        yb = fl_clamp(yb, 0.5f, hy);
        zb = fl_clamp(zb, 0.5f, hz);
        const float fi = floorf(xb), fj = floorf(yb), fk = floorf(zb);       // :392-397
        sx[j] = xb - fi; sy[j] = yb - fj; sz[j] = zb - fk;                   // :399-401
        const int i0 = (int)fi, j0 = (int)fj, k0 = (int)fk;
        if (SLAB) {
            const float* __restrict__ q0 = fl_gather_plane(g, src, k0) + (i0 + j0 * (int)pitch);
            const float* __restrict__ q1 = fl_gather_plane(g, src, k0 + 1) + (i0 + j0 * (int)pitch);
            q[j][0] = fl_ldnc(q0); q[j][1] = fl_ldnc(q0 + 1); q[j][2] = fl_ldnc(q0 + pitch); q[j][3] = fl_ldnc(q0 + pitch + 1);
            q[j][4] = fl_ldnc(q1); q[j][5] = fl_ldnc(q1 + 1); q[j][6] = fl_ldnc(q1 + pitch); q[j][7] = fl_ldnc(q1 + pitch + 1);
        } else {
            const unsigned o00 = FL_XOFF + (unsigned)i0 + (unsigned)j0 * pitch + (unsigned)(k0 + zoff) * plane;
            const unsigned o10 = o00 + pitch, o01 = o00 + plane, o11 = o01 + pitch;
            q[j][0] = fl_ldnc(gsrc + o00); q[j][1] = fl_ldnc(gsrc + o00 + 1); q[j][2] = fl_ldnc(gsrc + o10); q[j][3] = fl_ldnc(gsrc + o10 + 1);
            q[j][4] = fl_ldnc(gsrc + o01); q[j][5] = fl_ldnc(gsrc + o01 + 1); q[j][6] = fl_ldnc(gsrc + o11); q[j][7] = fl_ldnc(gsrc + o11 + 1);
        }
    }
    asm volatile("" ::: "memory");     // every gather is issued before the first store (the compiler would sink rows 1.. behind it)
#pragma unroll
    for (int j = 0; j < NY; ++j) {
        const int y = ya + j;
        if (y > g.H) break;
        const float tx = 1.0f - sx[j], ty = 1.0f - sy[j], tz = 1.0f - sz[j];
        const float c00 = q[j][0] * tx + q[j][1] * sx[j];                    // :412-415  (c000, c100)
        const float c01 = q[j][4] * tx + q[j][5] * sx[j];                    //           (c001, c101)
        const float c10 = q[j][2] * tx + q[j][3] * sx[j];                    //           (c010, c110)
        const float c11 = q[j][6] * tx + q[j][7] * sx[j];                    //           (c011, c111)
        const float e0 = c00 * ty + c10 * sy[j];                             // :417-418
        const float e1 = c01 * ty + c11 * sy[j];
        const float raw = (fl[j] & FL_SOLID) ? 0.0f : e0 * tz + e1 * sz[j];  // :420, :375-378
        // setBounds(b, field) (:183-246): ghosts take the pre-zero value, then solids / no-slip cells are zeroed
        const unsigned c = cc[j];
        field[c] = (fl[j] & zm) ? 0.0f : raw;
        if (xz_edge || y == 1 || y == g.H) {
            if (x == 1)   field[c - 1] = (b == 1) ? -raw : raw;              // :189
            if (x == g.W) field[c + 1] = raw;                                // :191
            const float gy = (b == 2) ? -raw : raw, gzv = (b == 3) ? -raw : raw;
            if (y == 1)   field[c - pitch] = gy;                             // :198-201
            if (y == g.H) field[c + pitch] = gy;
            if (gz == 1)   field[c - plane] = gzv;                           // :208-214
            if (gz == g.D) field[c + plane] = gzv;
        }
    }
}

// ---- the three velocity advections of step() (simulation.cpp:125-127) in ONE launch ------------------------
// advect(1, v_x, v_x_prev); advect(2, v_y, v_y_prev); advect(3, v_z, v_z_prev) read the velocity of the
// OTHER components only at their own cell (:380-382) and gather only from the *_prev copies, which none of
// them writes.  What call 2 reads of v_x at its own cell is what call 1 + setBounds(1) left there, and an
// interior cell's value after setBounds is its advected value, or zero where the zero mask of component 1
// says so -- computable in the thread.  So one thread produces v_x, then v_y from it, then v_z from both:
// one flag load, five own-cell loads instead of nine, one index set-up, three gathers, three stores,
// 32 instead of 48 B per cell.  Statements and operand order per component are those of k_advect_rows:
// bit-identical to the three launches (tests/test_gpu_parity.py).
struct GatherSrc3 { const float* base[3][8]; int zlo[8]; int dl[8]; int n; };

struct AdvConst {
    float kx, ky, kz, hx, hy, hz;
    unsigned pitch, plane;
    int zoff;
};

// back-trace from (fx, fy, fz) with velocity (vx, vy, vz), gather the 8 corners from component `comp`'s source
template <bool SLAB>
__device__ __forceinline__ void fl_adv_gather(const Grid& g, const GatherSrc3& src, int comp, const AdvConst& k,
                                              float fx, float fy, float fz, float vx, float vy, float vz,
                                              float& sx, float& sy, float& sz, float q[8])
{
    float xb = fx - k.kx * vx;                                               // :384-386
    float yb = fy - k.ky * vy;
    float zb = fz - k.kz * vz;
    xb = fl_clamp(xb, 0.5f, k.hx);                                           // :388-390
    yb = fl_clamp(yb, 0.5f, k.hy);
    zb = fl_clamp(zb, 0.5f, k.hz);
    const float fi = floorf(xb), fj = floorf(yb), fk = floorf(zb);           // :392-397
    sx = xb - fi; sy = yb - fj; sz = zb - fk;                                // :399-401
    const int i0 = (int)fi, j0 = (int)fj, k0 = (int)fk;
    if (SLAB) {
        int r0 = 0, r1 = 0;
#pragma unroll 1
        for (int t = 0; t < src.n - 1; ++t) {
            if (k0 >= src.zlo[t] + src.dl[t]) r0 = t + 1;
            if (k0 + 1 >= src.zlo[t] + src.dl[t]) r1 = t + 1;
        }
        const int o = FL_XOFF + i0 + j0 * (int)k.pitch;
        const float* __restrict__ q0 = src.base[comp][r0] + ((long long)(k0 - src.zlo[r0] + 1) * g.plane + o);
        const float* __restrict__ q1 = src.base[comp][r1] + ((long long)(k0 + 1 - src.zlo[r1] + 1) * g.plane + o);
        q[0] = fl_ldnc(q0); q[1] = fl_ldnc(q0 + 1); q[2] = fl_ldnc(q0 + k.pitch); q[3] = fl_ldnc(q0 + k.pitch + 1);
        q[4] = fl_ldnc(q1); q[5] = fl_ldnc(q1 + 1); q[6] = fl_ldnc(q1 + k.pitch); q[7] = fl_ldnc(q1 + k.pitch + 1);
    } else {
        const float* __restrict__ gs = src.base[comp][0];
        const unsigned o00 = FL_XOFF + (unsigned)i0 + (unsigned)j0 * k.pitch + (unsigned)(k0 + k.zoff) * k.plane;
        const unsigned o10 = o00 + k.pitch, o01 = o00 + k.plane, o11 = o01 + k.pitch;
        q[0] = fl_ldnc(gs + o00); q[1] = fl_ldnc(gs + o00 + 1); q[2] = fl_ldnc(gs + o10); q[3] = fl_ldnc(gs + o10 + 1);
        q[4] = fl_ldnc(gs + o01); q[5] = fl_ldnc(gs + o01 + 1); q[6] = fl_ldnc(gs + o11); q[7] = fl_ldnc(gs + o11 + 1);
    }
}

__device__ __forceinline__ float fl_adv_lerp(const float q[8], float sx, float sy, float sz)
{
    const float tx = 1.0f - sx, ty = 1.0f - sy, tz = 1.0f - sz;
    const float c00 = q[0] * tx + q[1] * sx;                                 // :412-415  (c000, c100)
    const float c01 = q[4] * tx + q[5] * sx;                                 //           (c001, c101)
    const float c10 = q[2] * tx + q[3] * sx;                                 //           (c010, c110)
    const float c11 = q[6] * tx + q[7] * sx;                                 //           (c011, c111)
    const float e0 = c00 * ty + c10 * sy;                                    // :417-418
    const float e1 = c01 * ty + c11 * sy;
    return e0 * tz + e1 * sz;                                                // :420
}

// the ghosts next to a cell on a domain face: setBounds(b, field) (:183-246) copies the cell's PRE-zero value
// (negated for the wall-normal component).  Out of line: one branch per row instead of 18 predicated stores.
__device__ __noinline__ void fl_adv_ghosts(int W, int H, int D, float* __restrict__ f1, float* __restrict__ f2,
                                           float* __restrict__ f3, unsigned c, float r1, float r2, float r3,
                                           int x, int y, int gz, unsigned pitch, unsigned plane)
{
    if (x == 1)   { f1[c - 1] = -r1; f2[c - 1] = r2; f3[c - 1] = r3; }       // :189
    if (x == W)   { f1[c + 1] = r1;  f2[c + 1] = r2; f3[c + 1] = r3; }       // :191
    if (y == 1)   { f1[c - pitch] = r1; f2[c - pitch] = -r2; f3[c - pitch] = r3; }   // :198-201
    if (y == H)   { f1[c + pitch] = r1; f2[c + pitch] = -r2; f3[c + pitch] = r3; }
    if (gz == 1)  { f1[c - plane] = r1; f2[c - plane] = r2; f3[c - plane] = -r3; }   // :208-214
    if (gz == D)  { f1[c + plane] = r1; f2[c + plane] = r2; f3[c + plane] = -r3; }
}

template <bool SLAB, int NY>
__global__ void __launch_bounds__(256) k_advect3(Grid g, float* __restrict__ vxf, float* __restrict__ vyf,
                                                 float* __restrict__ vzf, const float* __restrict__ vxp,
                                                 const float* __restrict__ vyp, const float* __restrict__ vzp,
                                                 GatherSrc3 src, const unsigned char* __restrict__ flags, float dt)
{
    const int x = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    const int ya = 1 + (blockIdx.y * blockDim.y + threadIdx.y) * NY;
    const int zl = 1 + blockIdx.z;
    if (x > g.W || ya > g.H) return;
    const int gz = g.zlo + zl - 1;
    const float fW = (float)g.W, fH = (float)g.H, fD = (float)g.D;
    AdvConst k;
    k.kx = dt * fW; k.ky = dt * fH; k.kz = dt * fD;                          // dt * width etc. (:384-386, left to right)
    k.hx = fW + 0.5f; k.hy = fH + 0.5f; k.hz = fD + 0.5f;
    k.pitch = (unsigned)g.pitch; k.plane = (unsigned)g.plane;
    k.zoff = 1 - src.zlo[0];
    const bool xz_edge = x == 1 || x == g.W || gz == 1 || gz == g.D;
    const float fx = (float)x, fz = (float)gz;
    const unsigned c0 = FL_XOFF + (unsigned)x + (unsigned)ya * k.pitch + (unsigned)zl * k.plane;
    // own-cell loads of every row first, then per component: all gathers of all rows in flight together
    unsigned cc[NY], fl[NY];
    float ox[NY], oy[NY], oz[NY], py[NY], pz[NY], fy[NY];
#pragma unroll
    for (int j = 0; j < NY; ++j) {
        const bool in = ya + j <= g.H;                                       // rows past H repeat row ya and are not stored
        cc[j] = in ? c0 + (unsigned)j * k.pitch : c0;
        fy[j] = (float)(in ? ya + j : ya);
        fl[j] = flags[cc[j]];
        ox[j] = vxp[cc[j]]; oy[j] = vyf[cc[j]]; oz[j] = vzf[cc[j]];          // advect(1): v_x_prev, v_y, v_z (:380-382)
        py[j] = vyp[cc[j]]; pz[j] = vzp[cc[j]];
    }
    float sx[NY], sy[NY], sz[NY], q[NY][8], r1[NY], r2[NY], r3[NY];
    // solid cells back-trace with zero velocity (their result is discarded, :375-378): the gather stays inside the volume
#pragma unroll
    for (int j = 0; j < NY; ++j) {
        const bool solid = fl[j] & FL_SOLID;
        fl_adv_gather<SLAB>(g, src, 0, k, fx, fy[j], fz, solid ? 0.0f : ox[j], solid ? 0.0f : oy[j], solid ? 0.0f : oz[j],
                            sx[j], sy[j], sz[j], q[j]);
    }
#pragma unroll
    for (int j = 0; j < NY; ++j) {
        r1[j] = (fl[j] & FL_SOLID) ? 0.0f : fl_adv_lerp(q[j], sx[j], sy[j], sz[j]);
        ox[j] = (fl[j] & fl_zero_mask(1)) ? 0.0f : r1[j];                    // v_x at this cell after advect(1) + setBounds(1)
    }
#pragma unroll
    for (int j = 0; j < NY; ++j) {                                           // advect(2): v_x (new), v_y_prev, v_z
        const bool solid = fl[j] & FL_SOLID;
        fl_adv_gather<SLAB>(g, src, 1, k, fx, fy[j], fz, solid ? 0.0f : ox[j], solid ? 0.0f : py[j], solid ? 0.0f : oz[j],
                            sx[j], sy[j], sz[j], q[j]);
    }
#pragma unroll
    for (int j = 0; j < NY; ++j) {
        r2[j] = (fl[j] & FL_SOLID) ? 0.0f : fl_adv_lerp(q[j], sx[j], sy[j], sz[j]);
        oy[j] = (fl[j] & fl_zero_mask(2)) ? 0.0f : r2[j];
    }
#pragma unroll
    for (int j = 0; j < NY; ++j) {                                           // advect(3): v_x (new), v_y (new), v_z_prev
        const bool solid = fl[j] & FL_SOLID;
        fl_adv_gather<SLAB>(g, src, 2, k, fx, fy[j], fz, solid ? 0.0f : ox[j], solid ? 0.0f : oy[j], solid ? 0.0f : pz[j],
                            sx[j], sy[j], sz[j], q[j]);
    }
    asm volatile("" ::: "memory");     // every gather is issued before the first store
#pragma unroll
    for (int j = 0; j < NY; ++j) {
        const int y = ya + j;
        if (y > g.H) break;
        r3[j] = (fl[j] & FL_SOLID) ? 0.0f : fl_adv_lerp(q[j], sx[j], sy[j], sz[j]);
        const unsigned c = cc[j];
        vxf[c] = (fl[j] & fl_zero_mask(1)) ? 0.0f : r1[j];
        vyf[c] = (fl[j] & fl_zero_mask(2)) ? 0.0f : r2[j];
        vzf[c] = (fl[j] & fl_zero_mask(3)) ? 0.0f : r3[j];
        if (xz_edge || y == 1 || y == g.H)
            fl_adv_ghosts(g.W, g.H, g.D, vxf, vyf, vzf, c, r1[j], r2[j], r3[j], x, y, gz, k.pitch, k.plane);
    }
}
