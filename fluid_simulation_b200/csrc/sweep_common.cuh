// sweep_common.cuh -- what the streaming sweep kernels share and what contains no PTX: tile constants, the
// segment record, 128-bit shared-memory accessors and the fused setBounds store.  Also compiled for the
// host by tests/host_emul/pair_emul.cpp.
#pragma once
#include "common.cuh"

#define SW_BOXW   136                    // floats per x box row: halo quad + 32 quads + halo quad
#define SW_RBOXW  128                    // floats per rhs box row (no halo columns)
#define SW_SMEM_MAX 232448               // 227 KB of dynamic shared memory per CTA on sm_100
#define SW_SM_SMEM  233472               // 228 KB of shared memory per SM (each resident CTA also reserves 1 KB)

#include "sweep_plan.h"                  // SwSeg and the work decomposition (pure host code)

// 128-bit shared load.  Where only two lanes of the quad are consumed the caller folds the other two
// into a sink word (sw_lds4_half), otherwise the load is narrowed to two scalar LDS.
__device__ __forceinline__ float4 sw_lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void   sw_sts4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// edge bits of a thread's quad (which face ghosts its store also has to write)
#define SW_XLO 1u
#define SW_XHI 2u
#define SW_YLO 4u
#define SW_YHI 8u
#define SW_ZLO 16u
#define SW_ZHI 32u

// the setBounds epilogue for one finished quad (simulation.cpp:183-246), see fl_store_bounded: ghosts take the
// PRE-zero value, then solids (and, for velocity components, cells touching a solid) are zeroed;
// bit i of `bits` = "cell i of the quad becomes 0"
__device__ __forceinline__ void sw_store_bits(const Grid& g, float* __restrict__ out, long long c, const float4& raw,
                                              unsigned bits, int b, unsigned edge)
{
    float4 v;
    v.x = (bits & 1u) ? 0.0f : raw.x;
    v.y = (bits & 2u) ? 0.0f : raw.y;
    v.z = (bits & 4u) ? 0.0f : raw.z;
    v.w = (bits & 8u) ? 0.0f : raw.w;
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
