// fluid_cuda.cu -- the C ABI of libfluid_cuda (include/fluid_cuda.h).  One translation unit:
// kernels live in the kernels_*.cuh / tiled.cuh / wavefront.cuh headers, orchestration in solver.cuh.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a --fmad=false -lineinfo (csrc/Makefile).
// There is no CPU path: without a usable CUDA device every entry point fails.
#include "solver.cuh"
#include "sweep_stream.cuh"
#include "sweep_pair.cuh"
#include "wavefront.cuh"
#include "comm.cuh"
#include "voxelize.cuh"
#include "postproc.cuh"
#include "mesh.cuh"
#include <cstring>
#include <new>
#include <unordered_map>
#include <mutex>

thread_local std::string g_fluid_create_error;

static Grid make_grid(int w, int h, int d, int zlo, int dl, int rank, int nranks)
{
    Grid g;
    g.W = w; g.H = h; g.D = d; g.Dl = dl; g.zlo = zlo;
    g.pitch = (w + 2 + FL_XOFF + 3) / 4 * 4;
    g.rows = h + 2;
    g.plane = (long long)g.pitch * g.rows;
    g.cells = g.plane * (dl + 2 + 2 * FL_ZPAD);
    g.zl_lo = rank == 0 ? 0 : 1 - FL_HALO;
    g.zl_hi = rank == nranks - 1 ? dl + 1 : dl + FL_HALO;
    g.nxq = (w + 3) / 4;
    return g;
}

static size_t ref_plane_floats(const Grid& g) { return (size_t)(g.W + 2) * g.rows; }

static int alloc_all(fluid_handle* h)
{
    const size_t fbytes = (size_t)h->g.cells * sizeof(float);
    // one zeroed guard row in front of and behind every field keeps the float4 / scalar
    // neighbour loads of ghost-adjacent quads inside the allocation
    // the handle's pointers address local plane 0; FL_ZPAD planes precede them
    for (int i = 0; i < FLUID_NFIELDS; ++i) {
        FL_CUDA(h, cudaMalloc(&h->field[i], fbytes));
        FL_CUDA(h, cudaMemsetAsync(h->field[i], 0, fbytes, h->stream));   // simulation.cpp:38-43
        h->alloc_base[i] = h->field[i];
        h->field[i] += FL_ZPAD * h->g.plane;
    }
    for (int i = 0; i < (h->nranks > 1 ? 4 : 3); ++i) {       // z-slabs: the fourth partner is IPC-mapped like the others
        FL_CUDA(h, cudaMalloc(&h->scratch[i], fbytes));
        FL_CUDA(h, cudaMemsetAsync(h->scratch[i], 0, fbytes, h->stream));
        h->alloc_base[FLUID_NFIELDS + i] = h->scratch[i];
        h->scratch[i] += FL_ZPAD * h->g.plane;
    }
    FL_CUDA(h, cudaMalloc(&h->cflags, (size_t)h->g.cells));
    FL_CUDA(h, cudaMemsetAsync(h->cflags, 0, (size_t)h->g.cells, h->stream));
    h->cflags += FL_ZPAD * h->g.plane;
    FL_CUDA(h, cudaMalloc(&h->staging, ref_plane_floats(h->g) * (h->g.Dl + 2 * FL_HALO) * sizeof(float)));
    h->n_partial = 148 * 8;
    FL_CUDA(h, cudaMalloc(&h->d_partial, (size_t)h->n_partial * FL_NSTAT * sizeof(double)));
    FL_CUDA(h, cudaMalloc(&h->d_out, (FL_NSTAT + 1) * sizeof(double)));        // + the statistics kernel's ticket
    FL_CUDA(h, cudaMemsetAsync(h->d_out, 0, (FL_NSTAT + 1) * sizeof(double), h->stream));
    FL_CUDA(h, cudaMalloc(&h->d_sync, 64 * sizeof(unsigned)));
    FL_CUDA(h, cudaMemsetAsync(h->d_sync, 0, 64 * sizeof(unsigned), h->stream));
    return 0;
}

static fluid_handle* create_impl(int w, int hgt, int d, int speed, float dt, float diff, float visc,
                                 int acc, unsigned flags, int rank, int nranks)
{
    if (w < 1 || hgt < 1 || d < 1 || acc < 0 || nranks < 1 || rank < 0 || rank >= nranks || nranks > 8 ||
        (nranks > 1 && d < FL_HALO * nranks)) {
        fl_fail(nullptr, FLUID_E_ARG, "fluid_create: bad arguments (w=%d h=%d d=%d acc=%d rank=%d/%d)",
                w, hgt, d, acc, rank, nranks);
        return nullptr;
    }
    if ((long long)(w + 2) * (hgt + 2) * (d + 2) > 2147483647LL) {
        fl_fail(nullptr, FLUID_E_ARG, "fluid_create: padded cell count exceeds int range (simulation.h:13)");
        return nullptr;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        fl_fail(nullptr, FLUID_E_CUDA, "fluid_create: no CUDA device (%s); libfluid_cuda has no CPU fallback",
                cudaGetErrorString(e));
        return nullptr;
    }
    fluid_handle* h = new (std::nothrow) fluid_handle();
    if (!h) { fl_fail(nullptr, FLUID_E_NOMEM, "fluid_create: out of host memory"); return nullptr; }
    cudaGetDevice(&h->device);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, h->device) == cudaSuccess) {
        h->sm_count = prop.multiProcessorCount;
        // L2 residency control for the sweep right-hand side (sweep_stream.cuh).  OFF by default:
        // measured on B200 at 256^3 the persisting carve-out takes so much L2 away from the streaming
        // x in/out traffic that a sweep slows from 45 to 100 us; FLUID_L2_PERSIST=1 enables it for A/B.
        const char* pe = getenv("FLUID_L2_PERSIST");
        if (pe && pe[0] == '1' && prop.persistingL2CacheMaxSize > 0 &&
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)prop.persistingL2CacheMaxSize) == cudaSuccess) {
            h->l2_persist_max = (size_t)prop.persistingL2CacheMaxSize;
            h->l2_window_max = (size_t)prop.accessPolicyMaxWindowSize;
        } else cudaGetLastError();
        if (prop.major < 10) {
            fl_fail(nullptr, FLUID_E_CUDA, "fluid_create: device sm_%d%d is not Blackwell (built for sm_100a only)",
                    prop.major, prop.minor);
            delete h; return nullptr;
        }
    }
    // equal z-slabs, remainder planes to the low ranks
    int base = d / nranks, rem = d % nranks;
    int dl = base + (rank < rem ? 1 : 0);
    int zlo = 1 + rank * base + (rank < rem ? rank : rem);
    h->g = make_grid(w, hgt, d, zlo, dl, rank, nranks);
    h->speed = speed; h->dt = dt; h->diff = diff; h->visc = visc; h->acc = acc; h->flags = flags;
    h->rank = rank; h->nranks = nranks;
    bool ok = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&h->dump_stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithPriority(&h->halo_stream, cudaStreamNonBlocking, -1) == cudaSuccess &&   // edge sweeps + halo exchange go first
              cudaEventCreateWithFlags(&h->ev_frame_packed, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&h->ev_frame_done, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&h->ev_a, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&h->ev_b, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreate(&h->ev_t0) == cudaSuccess && cudaEventCreate(&h->ev_t1) == cudaSuccess;
    if (!ok || alloc_all(h) != 0 || cudaStreamSynchronize(h->stream) != cudaSuccess) {
        std::string msg = h->err.empty() ? std::string("stream/event creation failed: ") +
                                           cudaGetErrorString(cudaGetLastError()) : h->err;
        fluid_destroy(h);
        fl_fail(nullptr, FLUID_E_CUDA, "fluid_create: %s", msg.c_str());
        return nullptr;
    }
    return h;
}

extern "C" {

const char* fluid_version(void)
{
    return "libfluid_cuda 0.1 (sm_100a, --fmad=false, built " __DATE__ ")";
}

fluid_handle* fluid_create(int w, int h, int d, int speed, float dt, float diff, float visc,
                           int acc, unsigned flags)
{
    return create_impl(w, h, d, speed, dt, diff, visc, acc, flags, 0, 1);
}

fluid_handle* fluid_create_slab(int w, int h, int d, int speed, float dt, float diff, float visc,
                                int acc, unsigned flags, int rank, int nranks)
{
    return create_impl(w, h, d, speed, dt, diff, visc, acc, flags, rank, nranks);
}

void fluid_destroy(fluid_handle* h)
{
    if (!h) return;
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->dump_stream) cudaStreamSynchronize(h->dump_stream);
    fl_tiled_release(h);
    fl_comm_release(h);
    for (int i = 0; i < FLUID_NFIELDS; ++i) if (h->field[i]) cudaFree(h->field[i] - FL_ZPAD * h->g.plane);
    for (int i = 0; i < 4; ++i) if (h->scratch[i]) cudaFree(h->scratch[i] - FL_ZPAD * h->g.plane);
    for (int i = 0; i < 5; ++i) cudaFree(h->frame_staging[i]);
    if (h->cflags) cudaFree(h->cflags - FL_ZPAD * h->g.plane);
    cudaFree(h->staging); cudaFree(h->d_partial); cudaFree(h->d_out);
    cudaFree(h->d_xyz); cudaFree(h->d_val); cudaFree(h->d_sync);
    for (auto& r : h->prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    for (auto e : h->ev_pool) cudaEventDestroy(e);
    if (h->ev_frame_packed) cudaEventDestroy(h->ev_frame_packed);
    if (h->ev_frame_done) cudaEventDestroy(h->ev_frame_done);
    if (h->ev_a) cudaEventDestroy(h->ev_a);
    if (h->ev_b) cudaEventDestroy(h->ev_b);
    if (h->ev_t0) cudaEventDestroy(h->ev_t0);
    if (h->ev_t1) cudaEventDestroy(h->ev_t1);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->dump_stream) cudaStreamDestroy(h->dump_stream);
    if (h->halo_stream) cudaStreamDestroy(h->halo_stream);
    delete h;
}

const char* fluid_last_error(fluid_handle* h) { return h ? h->err.c_str() : g_fluid_create_error.c_str(); }

int fluid_set_acc(fluid_handle* h, int acc)
{
    if (!h || acc < 0) return fl_fail(h, FLUID_E_ARG, "fluid_set_acc: bad argument");
    h->acc = acc;
    return 0;
}

int fluid_set_params(fluid_handle* h, int speed, float dt, float diff, float visc, int acc)
{
    if (!h || acc < 0) return fl_fail(h, FLUID_E_ARG, "fluid_set_params: bad argument");
    h->speed = speed; h->dt = dt; h->diff = diff; h->visc = visc; h->acc = acc;
    return 0;
}

int fluid_get_dims(fluid_handle* h, int* w, int* hh, int* d, int* z_lo, int* d_local)
{
    if (!h) return FLUID_E_ARG;
    if (w) *w = h->g.W; if (hh) *hh = h->g.H; if (d) *d = h->g.D;
    if (z_lo) *z_lo = h->g.zlo; if (d_local) *d_local = h->g.Dl;
    return 0;
}

int fluid_sync(fluid_handle* h)
{
    if (!h) return FLUID_E_ARG;
    FL_CUDA(h, cudaStreamSynchronize(h->stream));
    if (h->nranks > 1 && h->halo_seq > 0) {              // fused halo hand-over: did an arrival wait time out?
        unsigned err = 0;
        FL_CUDA(h, cudaMemcpy(&err, h->d_sync + 16, sizeof err, cudaMemcpyDeviceToHost));
        if (err) return fl_fail(h, FLUID_E_COMM, "z-neighbour %s never handed its halo planes over (arrival flag timed out)",
                                err == 1 ? "below" : "above");
    }
    return 0;
}

long long fluid_launch_count(fluid_handle* h) { return h ? h->launches : 0; }

int fluid_set_device(int device)
{
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fl_fail(nullptr, FLUID_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    return 0;
}

void* fluid_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void fluid_host_free(void* p) { if (p) cudaFreeHost(p); }

int fluid_timer_start(fluid_handle* h)
{
    if (!h) return FLUID_E_ARG;
    FL_CUDA(h, cudaEventRecord(h->ev_t0, h->stream));
    return 0;
}
int fluid_timer_stop(fluid_handle* h, float* ms)
{
    if (!h || !ms) return FLUID_E_ARG;
    FL_CUDA(h, cudaEventRecord(h->ev_t1, h->stream));
    FL_CUDA(h, cudaEventSynchronize(h->ev_t1));
    FL_CUDA(h, cudaEventElapsedTime(ms, h->ev_t0, h->ev_t1));
    return 0;
}

// per-solve device timing: every linearSolver call is bracketed by a CUDA event pair on the
// handle's stream while enabled; fluid_profile_read syncs, sums and resets.
int fluid_profile_enable(fluid_handle* h, int on)
{
    if (!h) return FLUID_E_ARG;
    h->profiling = on != 0;
    return 0;
}
int fluid_profile_read(fluid_handle* h, double out[3])
{
    if (!h || !out) return FLUID_E_ARG;
    FL_CUDA(h, cudaStreamSynchronize(h->stream));
    double ms = 0.0, sweeps = 0.0;
    for (auto& r : h->prof) {
        float t = 0.0f;
        FL_CUDA(h, cudaEventElapsedTime(&t, r.a, r.b));
        ms += t; sweeps += (double)r.field_sweeps;
        h->ev_pool.push_back(r.a); h->ev_pool.push_back(r.b);
    }
    out[0] = ms; out[1] = sweeps; out[2] = (double)h->prof.size();
    h->prof.clear();
    return 0;
}

}  // extern "C"

#include "api_io.inl"
#include "api_ops.inl"

// ---- fluid_stats_static: run()'s statistics without any allocation ------------------------------------
// No counterpart in the reference; added for callers that cannot
// call cudaMalloc (CUDA graph capture, tight serving loops).  A two-pass reduction with fluid_stats' block
// fold, but the per-block partials live in a file-scope device array instead of the handle's cudaMalloc'ed
// d_partial, and the fields are the caller's flat device arrays.
#define FL_STATIC_CHUNK      65536                  // cells folded by one block
#define FL_STATIC_MAX_BLOCKS 4096                   // => at most 2^28 cells per call
#define FL_STATIC_MAX_DEVICES 64
// 4096 x 9 doubles = 288 KiB, held by every GPU that loads the library until the process exits
static __device__ __align__(256) double g_static_partials[FL_STATIC_MAX_BLOCKS * FL_NSTAT];
// the array's DEVICE address per device (the name above is only a host-side stand-in in host code)
static double* g_static_partials_addr[FL_STATIC_MAX_DEVICES];
static std::mutex g_static_mutex;

// block b folds cells [b*CHUNK, min(n, (b+1)*CHUNK)) of the four flat arrays
__global__ void __launch_bounds__(256) k_stats_flat_partial(long long n, const float* __restrict__ dens,
                                                            const float* __restrict__ vx,
                                                            const float* __restrict__ vy,
                                                            const float* __restrict__ vz,
                                                            double* __restrict__ partial)
{
    const long long lo = (long long)blockIdx.x * FL_STATIC_CHUNK;
    const long long hi = lo + FL_STATIC_CHUNK < n ? lo + FL_STATIC_CHUNK : n;
    double s = 0.0;
    float mn[4] = { FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX }, mx[4] = { -FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (long long c = lo + threadIdx.x; c < hi; c += blockDim.x) {
        const float d = dens[c], a = vx[c], b = vy[c], e = vz[c];
        s += (double)d;
        mn[0] = fminf(mn[0], d); mx[0] = fmaxf(mx[0], d);
        mn[1] = fminf(mn[1], a); mx[1] = fmaxf(mx[1], a);
        mn[2] = fminf(mn[2], b); mx[2] = fmaxf(mx[2], b);
        mn[3] = fminf(mn[3], e); mx[3] = fmaxf(mx[3], e);
    }
    block_fold_stats(s, mn, mx, partial + (long long)blockIdx.x * FL_NSTAT);
}

// set-up, once per device, with that device current: the device address of g_static_partials
static int fl_static_scratch(double** out)
{
    int dev = -1;
    FL_CUDA(nullptr, cudaGetDevice(&dev));
    if (dev < 0 || dev >= FL_STATIC_MAX_DEVICES)
        return fl_fail(nullptr, FLUID_E_ARG, "fluid_stats_static: device ordinal %d outside 0..%d", dev,
                       FL_STATIC_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_partials_addr[dev]) {
        double* p = nullptr;
        FL_CUDA(nullptr, cudaGetSymbolAddress((void**)&p, g_static_partials));
        if (!p) return fl_fail(nullptr, FLUID_E_CUDA, "fluid_stats_static: cudaGetSymbolAddress returned NULL");
        g_static_partials_addr[dev] = p;
    }
    *out = g_static_partials_addr[dev];
    return 0;
}

extern "C" int fluid_stats_static(const float* dens, const float* v_x, const float* v_y, const float* v_z,
                                  long long n_cells, double* out_device, void* stream)
{
    if (!dens || !v_x || !v_y || !v_z || !out_device)
        return fl_fail(nullptr, FLUID_E_ARG, "fluid_stats_static: NULL field or output pointer");
    const long long max_cells = (long long)FL_STATIC_MAX_BLOCKS * FL_STATIC_CHUNK;
    if (n_cells < 1 || n_cells > max_cells)
        return fl_fail(nullptr, FLUID_E_ARG,
                       "fluid_stats_static: n_cells = %lld outside 1..%lld (the static scratch holds %d block "
                       "partials of %d cells each; use fluid_stats for larger volumes)",
                       n_cells, max_cells, FL_STATIC_MAX_BLOCKS, FL_STATIC_CHUNK);
    double* partial = nullptr;
    if (int e = fl_static_scratch(&partial)) return e;
    cudaStream_t s = (cudaStream_t)stream;
    const int nb = (int)((n_cells + FL_STATIC_CHUNK - 1) / FL_STATIC_CHUNK);
    // every partial the second pass reads is written by the first pass of the SAME call: nothing in the
    // shared array has to start at zero, and what the previous call left there is never read
    k_stats_flat_partial<<<nb, 256, 0, s>>>(n_cells, dens, v_x, v_y, v_z, partial);
    FL_CUDA(nullptr, cudaGetLastError());
    k_stats_final<<<1, 256, 0, s>>>(nb, partial, out_device);
    FL_CUDA(nullptr, cudaGetLastError());
    return 0;
}
