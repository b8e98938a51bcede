// recip_div.h -- x / s for a divisor s that is the same for a whole launch (project()'s h and 2h,
// simulation.cpp:330-334), bit-identical to the IEEE division, without the division.
//
// With r = RN(1/s):  q0 = RN(x*r);  e = x - q0*s (one FMA; exact, because q0 is within 2 ulp of x/s);
// q = RN(q0 + e*r) (one FMA).  That q equals RN(x/s) is not taken on trust: scaling x by a power of two
// scales q0, e and q exactly (as long as nothing leaves the normal range -- the guard below), and the
// rounding is symmetric in the sign, so the statement only has to hold for the 2^23 significands of one
// binade.  fl_recip_div_verify() checks exactly those, on the host, with the same three operations, once
// per divisor and process (~40 ms); a divisor that fails -- none has been seen -- keeps the division.
// |x| outside [2^-60, 2^60] (zeros, denormals, infinities, NaNs) and divisors outside [2^-24, 2^8] take the
// division too; a zero keeps its sign without one.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define FL_RD_FN __host__ __device__ __forceinline__
#else
#define FL_RD_FN static inline
#endif

FL_RD_FN float fl_recip_div_core(float x, float s, float r)
{
#ifdef __CUDA_ARCH__
    const float q0 = __fmul_rn(x, r);
    const float e = __fmaf_rn(-q0, s, x);
    return __fmaf_rn(e, r, q0);
#else
    const float q0 = x * r;
    const float e = __builtin_fmaf(-q0, s, x);
    return __builtin_fmaf(e, r, q0);
#endif
}

// the rare operands: out of line, so that the common path is three instructions and a branch, not the
// division's own fast and slow paths evaluated beside it
#ifdef __CUDACC__
static __host__ __device__ __noinline__
#else
static __attribute__((noinline))
#endif
float fl_recip_div_rare(float x, float s)
{
    return x / s;
}

FL_RD_FN float fl_recip_div(float x, float s, float r)
{
    const float ax = fabsf(x);
    if (ax >= 0x1p-60f && ax <= 0x1p60f) return fl_recip_div_core(x, s, r);
    if (x == 0.0f) return x;                           // (+-0) / s = +-0 for s > 0
    return fl_recip_div_rare(x, s);
}

// host: is fl_recip_div_core(x, s, 1/s) == x / s for every x in the guarded range?
#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target("fma")))
#endif
static bool fl_recip_div_sweep(float s, float r)
{
    for (unsigned m = 0; m < (1u << 23); ++m) {
        union { unsigned u; float f; } x;
        x.u = 0x3f800000u | m;
        const float q0 = x.f * r;
        const float e = __builtin_fmaf(-q0, s, x.f);
        if (__builtin_fmaf(e, r, q0) != x.f / s) return false;
    }
    return true;
}
static inline bool fl_recip_div_verify(float s)
{
    if (!(s >= 0x1p-24f && s <= 0x1p8f)) return false;
#if defined(__x86_64__) && defined(__GNUC__)
    if (!__builtin_cpu_supports("fma")) return false;          // the sweep is compiled for hardware FMA
#endif
    return fl_recip_div_sweep(s, 1.0f / s);
}
