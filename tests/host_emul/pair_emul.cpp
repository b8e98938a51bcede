// pair_emul.cpp -- TEST INFRASTRUCTURE.  Compiles the device code of the temporally blocked sweep
// (fluid_simulation_b200/csrc/sweep_pair_kernel.cuh) for the HOST and runs it CTA by CTA with one host
// thread per CUDA thread: __syncthreads is a std::barrier, warp shuffles go through a per-warp exchange
// buffer, the TMA box copies (zero fill outside the array, transaction bytes counted against the
// mbarrier) are done by the issuing thread.  tests/test_pair_emul.py compares the result bit for bit with
// the ordering-matched CPU oracle, so the kernel's statements are checked where no GPU is at hand.
// Built by the test with: g++ -std=c++20 -O1 -ffp-contract=off -shared -fPIC -pthread -I<cuda>/include
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>
#include <algorithm>

#define SP_HOST_EMUL 1
#include <cuda_runtime.h>                 // float4, make_float4, the (ignored) function qualifiers
using std::max;
using std::min;
#ifndef __noinline__
#define __noinline__ __attribute__((noinline))
#endif
static thread_local uint3 threadIdx, blockIdx;           // only named by helpers the emulation never calls
static thread_local dim3 blockDim, gridDim;

// ---- what the device code expects from the CUDA environment -------------------------------------
struct CUtensorMap {                      // 3-D float tensor, box dims[0] x dims[1] x 1
    const float* base; long long dims[3]; long long strides[2] /* floats */; int box[2];
};
struct EmulCta {
    int nthreads;
    std::barrier<> cta_bar;
    std::vector<std::unique_ptr<std::barrier<>>> warp_bar;
    std::vector<float> wbuf;
    explicit EmulCta(int n) : nthreads(n), cta_bar(n), wbuf((size_t)n) {
        for (int i = 0; i < n / 32; ++i) warp_bar.emplace_back(new std::barrier<>(32));
    }
};
static thread_local EmulCta* t_cta = nullptr;
static thread_local int t_tid = 0;

static inline unsigned __float_as_uint(float f) { unsigned u; memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(unsigned u) { float f; memcpy(&f, &u, 4); return f; }
static inline void __syncthreads() { t_cta->cta_bar.arrive_and_wait(); }
static inline float emul_shfl(float v, int src_delta)
{
    const int warp = t_tid >> 5, lane = t_tid & 31;
    t_cta->wbuf[t_tid] = v;
    t_cta->warp_bar[warp]->arrive_and_wait();
    const int src = lane + src_delta;
    const float r = (src >= 0 && src < 32) ? t_cta->wbuf[(warp << 5) + src] : v;
    t_cta->warp_bar[warp]->arrive_and_wait();
    return r;
}
static inline float __shfl_up_sync(unsigned, float v, int d) { return emul_shfl(v, -d); }
static inline float __shfl_down_sync(unsigned, float v, int d) { return emul_shfl(v, d); }

// mbarrier: low 32 bits = completed phases, high 32 bits = pending transaction bytes
static inline void sw_mbar_init(unsigned long long* bar, unsigned) { __atomic_store_n(bar, 0ull, __ATOMIC_SEQ_CST); }
static inline void sw_fence_mbar_init() {}
static inline void sw_grid_dep_sync() {}
static inline void sw_mbar_expect_tx(unsigned long long* bar, unsigned bytes)
{
    __atomic_fetch_add(bar, (unsigned long long)bytes << 32, __ATOMIC_SEQ_CST);
}
static inline void sw_mbar_wait(unsigned long long* bar, unsigned parity)
{
    for (;;) {
        const unsigned long long v = __atomic_load_n(bar, __ATOMIC_SEQ_CST);
        if (((unsigned)v & 1u) != parity) return;
        std::this_thread::yield();
    }
}
static inline void sw_tma_load_3d(void* dst, const CUtensorMap* m, int c0, int c1, int c2, unsigned long long* bar)
{
    float* d = (float*)dst;
    for (int r = 0; r < m->box[1]; ++r)
        for (int i = 0; i < m->box[0]; ++i) {
            const long long x = c0 + i, y = c1 + r, z = c2;
            const bool in = x >= 0 && x < m->dims[0] && y >= 0 && y < m->dims[1] && z >= 0 && z < m->dims[2];
            d[(size_t)r * m->box[0] + i] = in ? m->base[x + y * m->strides[0] + z * m->strides[1]] : 0.0f;
        }
    const unsigned long long bytes = (unsigned long long)m->box[0] * m->box[1] * 4ull;
    unsigned long long v = __atomic_sub_fetch(bar, bytes << 32, __ATOMIC_SEQ_CST);
    if ((v >> 32) == 0) __atomic_fetch_add(bar, 1ull, __ATOMIC_SEQ_CST);     // all bytes arrived: phase completes
}

#include "../../fluid_simulation_b200/csrc/sweep_pair_kernel.cuh"

// ---- harness --------------------------------------------------------------------------------------
static Grid make_grid(int w, int h, int d)
{
    Grid g;
    g.W = w; g.H = h; g.D = d; g.Dl = d; g.zlo = 1;
    g.pitch = (w + 2 + FL_XOFF + 3) / 4 * 4;
    g.rows = h + 2;
    g.plane = (long long)g.pitch * g.rows;
    g.cells = g.plane * (d + 2 + 2 * FL_ZPAD);
    g.zl_lo = 0; g.zl_hi = d + 1;
    g.nxq = (w + 3) / 4;
    return g;
}

template <bool UP, int NU>
static void run_cta(const PairArgs& A, int bid)
{
    using GE = SpGeom<8>;
    EmulCta cta(GE::THREADS);
    std::vector<float> smem(GE::SMEM / sizeof(float) + 64, std::nanf(""));   // uninitialised shared memory = NaN
    float* sm = (float*)(((uintptr_t)smem.data() + 127) & ~(uintptr_t)127);
    unsigned long long bars[SP_NR];
    std::vector<std::thread> th;
    for (int t = 0; t < GE::THREADS; ++t)
        th.emplace_back([&, t]() { t_cta = &cta; t_tid = t; sp_cta<8, UP, NU>(A, sm, bars, t, bid); });
    for (auto& x : th) x.join();
}

// One launch of k_sweep_pair on a single field: x, rhs, obs, out are padded volumes in the reference
// layout ((d+2) x (h+2) x (w+2)); segments are columns cut into chunks of `zchunk` planes, two segments per
// CTA where possible (so that the ring runs on across segments).  Returns 0, or a negative error.
extern "C" int pair_emul_sweep(int w, int h, int d, const float* x, const float* rhs, const float* obs, int b,
                               float a, float c, int zchunk, int up, int unroll, float* out)
{
    using GE = SpGeom<8>;
    if (w <= 128 || w > 256 || (w & 7) || (h & 1)) return -1;
    const Grid g = make_grid(w, h, d);
    const size_t n = (size_t)g.cells;
    std::vector<float> fx(n, 0.0f), fr(n, 0.0f), fo(n, 0.0f);
    std::vector<unsigned char> fl(n, 0);
    const int Wp = w + 2;
    auto unpack = [&](const float* src, std::vector<float>& dst) {
        for (int z = 0; z <= d + 1; ++z) for (int y = 0; y <= h + 1; ++y) for (int xx = 0; xx <= w + 1; ++xx)
            dst[(size_t)(FL_ZPAD * g.plane) + fl_idx(g, xx, y, z)] = src[((size_t)z * g.rows + y) * Wp + xx];
    };
    unpack(x, fx); unpack(rhs, fr);
    fo = fx;                                               // the partner buffer starts as a copy (edges, corners)
    // zero bits as k_build_zbits makes them (solid: obs == 1; no-slip: an in-domain 6-neighbour is solid)
    const int pairs = (g.rows + 1) / 2;
    std::vector<unsigned short> z0((size_t)(d + 2 * FL_HALO) * pairs * 32, 0), z1 = z0;
    auto ob = [&](int xx, int y, int z) { return obs[((size_t)z * g.rows + y) * Wp + xx]; };
    for (int z = 1; z <= d; ++z) for (int y = 1; y <= h; ++y) for (int xx = 1; xx <= w; ++xx) {
        const bool solid = ob(xx, y, z) == 1.0f;
        const bool ns = !solid && ((xx + 1 <= w && ob(xx + 1, y, z) == 1.0f) || (xx - 1 >= 1 && ob(xx - 1, y, z) == 1.0f) ||
                                   (y + 1 <= h && ob(xx, y + 1, z) == 1.0f) || (y - 1 >= 1 && ob(xx, y - 1, z) == 1.0f) ||
                                   (z + 1 <= d && ob(xx, y, z + 1) == 1.0f) || (z - 1 >= 1 && ob(xx, y, z - 1) == 1.0f));
        const size_t t = ((size_t)(z + FL_HALO - 1) * pairs + (y >> 1)) * 32 + (xx - 1) / 8;
        const unsigned bit = 1u << (8 * (y & 1) + ((xx - 1) & 7));
        if (solid) z0[t] |= bit;
        if (solid || ns) z1[t] |= bit;
    }
    // segments
    const int nty = (h + GE::TY - 1) / GE::TY;
    std::vector<SwSeg> segs; std::vector<int> first;
    int k = 0;
    for (int col = 0; col < nty; ++col)
        for (int za = 1; za <= d; za += zchunk, ++k) {
            if ((k & 1) == 0) first.push_back((int)segs.size());
            segs.push_back(SwSeg{ col, za, std::min(za + zchunk - 1, d), 0 });
        }
    first.push_back((int)segs.size());

    PairArgs A;
    memset(&A, 0, sizeof A);
    A.g = g; A.a = a; A.c_recip = 1.0f / c; A.nty = nty;
    A.segs = segs.data(); A.seg_first = first.data();
    A.zbits[0] = z0.data(); A.zbits[1] = z1.data(); A.zb_pairs = pairs;
    A.b[0] = b; A.dst[0] = fo.data() + FL_ZPAD * g.plane;
    A.map[0] = CUtensorMap{ fx.data(), { g.pitch, g.rows, g.Dl + 2 + 2 * FL_ZPAD }, { g.pitch, g.plane }, { SW_BOXW, GE::BR } };
    A.rmap[0] = CUtensorMap{ fr.data(), { g.pitch, g.rows, g.Dl + 2 + 2 * FL_ZPAD }, { g.pitch, g.plane }, { SW_RBOXW, GE::TR } };
    for (int bid = 0; bid + 1 < (int)first.size(); ++bid) {
        if (unroll == 6) { if (up) run_cta<true, 6>(A, bid); else run_cta<false, 6>(A, bid); }
        else             { if (up) run_cta<true, 2>(A, bid); else run_cta<false, 2>(A, bid); }
    }
    for (int z = 0; z <= d + 1; ++z) for (int y = 0; y <= h + 1; ++y) for (int xx = 0; xx <= w + 1; ++xx)
        out[((size_t)z * g.rows + y) * Wp + xx] = fo[(size_t)(FL_ZPAD * g.plane) + fl_idx(g, xx, y, z)];
    return 0;
}
