// handle.cuh -- the state behind `fluid_handle*` and launch helpers.
#pragma once
#include "common.cuh"
#include "../../include/fluid_cuda.h"
#include <string>
#include <vector>
#include <cstdio>
#include <cstdarg>

struct fluid_handle {
    Grid g;
    int speed; float dt, diff, visc; int acc; unsigned flags;
    int rank, nranks;
    int device;
    cudaStream_t stream = nullptr, dump_stream = nullptr, halo_stream = nullptr;
    cudaEvent_t ev_frame_packed = nullptr, ev_frame_done = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
    cudaEvent_t ev_a = nullptr, ev_b = nullptr;
    float* field[FLUID_NFIELDS] = {};
    float* scratch[4] = {};          // ping-pong partners of the batched solves ([3]: z-slabs at creation, one GPU on first use)
    // Allocation starts (plane -1) in FLUID_IPC_* id order.  field[] / scratch[] are permuted by the
    // solver's pointer swaps, identically on every rank; a buffer's id is found by its address.
    float* alloc_base[FLUID_NFIELDS + 4] = {};
    unsigned* d_sync = nullptr;      // [0] / [1] = arrival flags raised by the lower / upper z-neighbour, [2] = CTA counter
    unsigned halo_seq = 0;           // halo hand-overs (sweeps, exchanges) that went out through peer stores so far
    unsigned barrier_seq = 0;        // flag barriers so far (d_sync[32 + r] = epoch of rank r)
    unsigned char* cflags = nullptr; // per-cell flag bytes, same pitched geometry as the fields
    bool flags_dirty = true;
    unsigned long long flags_version = 0;   // bumped whenever the flag volume is rebuilt (derived tables follow it)
    float* staging = nullptr;        // (Dl+2)*(H+2)*(W+2) floats, reference layout, for up/download
    float* frame_staging[5] = {};    // lazily allocated, reference layout, for async frame dumps
    bool frame_in_flight = false;
    double* d_partial = nullptr; double* d_out = nullptr; int n_partial = 0;
    int* d_xyz = nullptr; float* d_val = nullptr; int sparse_cap = 0;
    long long launches = 0;
    std::string err;
    // multi-GPU halo exchange hook (slab mode); see fluid_cuda.cu
    void* comm = nullptr;
    void* tiled = nullptr;           // TiledState (tensor maps), tiled.cuh
    // optional per-solve CUDA-event timing (fluid_profile_*): one event pair per linearSolver call
    bool profiling = false;
    struct ProfRec { cudaEvent_t a, b; long long field_sweeps; };
    std::vector<ProfRec> prof;
    std::vector<cudaEvent_t> ev_pool;
    // set by fl_step_only when it swapped v <-> v_prev instead of copying: the next tiled solve takes its initial
    // guess from the right-hand-side buffer (which holds the very same values) and clears the flag
    unsigned x_from_rhs = 0;         // bit i: batched field i
    int slab_swap_ok = -1;           // z-slabs: v <-> v_prev swap agreed by all ranks (-1: not asked yet)
    int sm_count = 148;
    size_t l2_persist_max = 0, l2_window_max = 0;   // persisting-L2 carve-out and access-window limits
};

extern thread_local std::string g_fluid_create_error;

inline int fl_fail(fluid_handle* h, int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (h) h->err = buf; else g_fluid_create_error = buf;
    return code;
}

#define FL_CUDA(h, call)                                                                        \
    do { cudaError_t e_ = (call);                                                               \
         if (e_ != cudaSuccess)                                                                 \
             return fl_fail((h), FLUID_E_CUDA, "%s failed: %s (%s:%d)", #call,                  \
                            cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

#define FL_CHECK_LAUNCH(h)                                                                      \
    do { cudaError_t e_ = cudaGetLastError();                                                   \
         if (e_ != cudaSuccess)                                                                 \
             return fl_fail((h), FLUID_E_CUDA, "kernel launch failed: %s (%s:%d)",              \
                            cudaGetErrorString(e_), __FILE__, __LINE__);                        \
         (h)->launches++; } while (0)

struct QuadLaunch { dim3 grid, block; };

// blockDim = (bx, by) with bx = smallest power of two covering the x-quads (<= 64), 256 threads
inline QuadLaunch fl_quad_launch(const Grid& g, int planes)
{
    int bx = 1;
    while (bx < g.nxq && bx < 64) bx <<= 1;
    int by = 256 / bx;
    QuadLaunch q;
    q.block = dim3(bx, by, 1);
    q.grid = dim3((g.nxq + bx - 1) / bx, (g.H + by - 1) / by, planes);
    return q;
}

inline int fl_grid_1d(long long n, int block = 256, int cap = 148 * 16)
{
    long long b = (n + block - 1) / block;
    if (b < 1) b = 1;
    if (b > cap) b = cap;
    return (int)b;
}
