// comm.cuh -- multi-GPU z-slab plumbing: one process per GPU, rank r owns a contiguous
// z-slab.  Two mechanisms, both set up by the host program (fluid_simulation_b200/multi_gpu.py):
//   * NCCL (ncclSend / ncclRecv in one group per exchange, on the solver stream) moves the two halo
//     planes per side after every sweep and the one-off exchanges around divergence / gradient / advect.
//     A plane is one contiguous run of pitch * (H+2) floats, so nothing is packed.  libnccl is dlopen'ed
//     (the torch-bundled library of the host process), the communicator is the library's own.
//   * CUDA IPC maps the other ranks' advection sources (v_*_prev, buffer) into this process: the
//     back-traces of advect() reach arbitrarily far in z, so k_advect gathers straight from
//     the peer slab over NVLink instead of exchanging a halo of unknown depth.
#pragma once
#include "handle.cuh"
#include <nccl.h>
#include <cuda.h>
#include <dlfcn.h>

struct SlabComm {
    void* lib = nullptr;
    ncclComm_t comm = nullptr;
    float* d_token = nullptr;                              // one float: all-reduce used as a device-side barrier
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    // IPC-mapped allocations of the other ranks, by allocation slot (fluid_handle::alloc_base order;
    // [FLUID_IPC_SYNC] = the rank's flag block)
    float* peer_alloc[8][FLUID_IPC_COUNT] = {};
    // stream memory operation (driver entry point): the stream waits until *addr >= value
    CUresult (*WaitValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned) = nullptr;
    int push_state = 0;                                    // 0 = not probed yet, 1 = peer stores usable, -1 = NCCL only
};

static int comm_load(SlabComm* c, const char* path, std::string* err)
{
    if (c->lib) return 0;
    const char* cand[3] = { path && path[0] ? path : nullptr, "libnccl.so.2", "libnccl.so" };
    for (int i = 0; i < 3 && !c->lib; ++i) if (cand[i]) c->lib = dlopen(cand[i], RTLD_NOW | RTLD_GLOBAL);
    if (!c->lib) { *err = std::string("cannot dlopen libnccl: ") + dlerror(); return FLUID_E_COMM; }
#define FL_SYM(field, name) *(void**)(&c->field) = dlsym(c->lib, name); if (!c->field) { *err = "libnccl lacks " name; return FLUID_E_COMM; }
    FL_SYM(GetUniqueId, "ncclGetUniqueId") FL_SYM(CommInitRank, "ncclCommInitRank") FL_SYM(CommDestroy, "ncclCommDestroy")
    FL_SYM(Send, "ncclSend") FL_SYM(Recv, "ncclRecv") FL_SYM(AllReduce, "ncclAllReduce")
    FL_SYM(GroupStart, "ncclGroupStart") FL_SYM(GroupEnd, "ncclGroupEnd") FL_SYM(GetErrorString, "ncclGetErrorString")
#undef FL_SYM
    return 0;
}

#define FL_NCCL(h, c, call)                                                                        \
    do { ncclResult_t r_ = (call);                                                                 \
         if (r_ != ncclSuccess)                                                                    \
             return fl_fail((h), FLUID_E_COMM, "%s failed: %s (%s:%d)", #call,                     \
                            (c)->GetErrorString(r_), __FILE__, __LINE__); } while (0)

// equal z-slabs, remainder planes to the low ranks: the rule of create_impl / slab.partition
static inline void fl_slab_of(int d, int nranks, int r, int* zlo, int* dl)
{
    const int base = d / nranks, rem = d % nranks;
    *dl = base + (r < rem ? 1 : 0);
    *zlo = 1 + r * base + (r < rem ? r : rem);
}

// Send my planes 1 .. FL_HALO down and Dl-FL_HALO+1 .. Dl up, receive the neighbours' into Dl+1 .. Dl+FL_HALO /
// 1-FL_HALO .. 0 (every slab has at least FL_HALO planes).
int fl_halo_exchange_on(fluid_handle* h, int n, float* const* fields, cudaStream_t stream)
{
    if (h->nranks == 1) return 0;
    SlabComm* c = (SlabComm*)h->comm;
    if (!c || !c->comm) return fl_fail(h, FLUID_E_COMM, "halo exchange: no communicator attached to this slab handle (fluid_comm_init)");
    const size_t cnt = FL_HALO * (size_t)h->g.plane;
    const long long pl = h->g.plane;
    const int Dl = h->g.Dl;
    FL_NCCL(h, c, c->GroupStart());
    for (int i = 0; i < n; ++i) {
        float* f = fields[i];
        if (h->rank > 0) {
            FL_NCCL(h, c, c->Send(f + 1 * pl, cnt, ncclFloat, h->rank - 1, c->comm, stream));
            FL_NCCL(h, c, c->Recv(f - (long long)(FL_HALO - 1) * pl, cnt, ncclFloat, h->rank - 1, c->comm, stream));
        }
        if (h->rank < h->nranks - 1) {
            FL_NCCL(h, c, c->Send(f + (long long)(Dl - FL_HALO + 1) * pl, cnt, ncclFloat, h->rank + 1, c->comm, stream));
            FL_NCCL(h, c, c->Recv(f + (long long)(Dl + 1) * pl, cnt, ncclFloat, h->rank + 1, c->comm, stream));
        }
    }
    FL_NCCL(h, c, c->GroupEnd());
    h->launches++;
    return 0;
}

// The exchange around divergence / gradient / advect.  With the neighbours' buffers mapped it is a hand-over like
// the sweeps': one small kernel stores the FL_HALO boundary planes per side into the neighbours' halos and raises
// their arrival flags, the stream then waits for its own two flags -- no rendezvous with NCCL's proxy on the
// step's critical path.  One-sided stores are safe here because between any kernel that READS a halo and the
// next hand-over INTO that halo every rank passes at least one other hand-over whose flag the writer waited for.
int fl_halo_push(fluid_handle* h, int n, float* const* fields, cudaStream_t stream, unsigned seq, int nplanes);
int fl_halo_wait(fluid_handle* h, cudaStream_t stream, unsigned seq);
bool fl_push_usable(fluid_handle* h);
int fl_halo_exchange(fluid_handle* h, int n, float* const* fields)
{
    if (h->nranks == 1) return 0;
    static const bool nccl_only = getenv("FLUID_XCHG_NCCL") != nullptr;   // A/B: ncclSend / ncclRecv as before
    if (!nccl_only && fl_push_usable(h)) {
        const unsigned seq = ++h->halo_seq;
        if (int e = fl_halo_push(h, n, fields, h->stream, seq, FL_HALO)) return e;
        return fl_halo_wait(h, h->stream, seq);
    }
    return fl_halo_exchange_on(h, n, fields, h->stream);
}

struct BarrierArgs { unsigned* flag[8]; unsigned epoch; int nranks, rank; };
// every earlier kernel of this stream is complete (stream order); fence, then tell every other rank
__global__ void k_barrier_signal(const BarrierArgs a)
{
    __threadfence_system();
    const int r = threadIdx.x;
    if (r < a.nranks && r != a.rank && a.flag[r])
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.flag[r] + 32 + a.rank), "r"(a.epoch) : "memory");
}

// all ranks' streams have reached this point before any continues: epoch flags over NVLink where every
// rank's flag block is mapped, else a 1-float all-reduce
static inline bool fl_push_ready(fluid_handle* h);
int fl_barrier(fluid_handle* h)
{
    if (h->nranks == 1) return 0;
    SlabComm* c = (SlabComm*)h->comm;
    if (!c) return fl_fail(h, FLUID_E_COMM, "barrier: no communicator attached");
    static const bool nccl_only = getenv("FLUID_XCHG_NCCL") != nullptr;
    bool all_mapped = !nccl_only && fl_push_ready(h);
    for (int r = 0; r < h->nranks && all_mapped; ++r) if (r != h->rank && !c->peer_alloc[r][FLUID_IPC_SYNC]) all_mapped = false;
    if (all_mapped) {
        BarrierArgs a = {};
        a.epoch = ++h->barrier_seq; a.nranks = h->nranks; a.rank = h->rank;
        for (int r = 0; r < h->nranks; ++r) a.flag[r] = r == h->rank ? nullptr : (unsigned*)c->peer_alloc[r][FLUID_IPC_SYNC];
        k_barrier_signal<<<1, 32, 0, h->stream>>>(a);
        FL_CHECK_LAUNCH(h);
        for (int r = 0; r < h->nranks; ++r) {
            if (r == h->rank) continue;
            CUresult e = c->WaitValue32((CUstream)h->stream, (CUdeviceptr)(h->d_sync + 32 + r), a.epoch, CU_STREAM_WAIT_VALUE_GEQ);
            if (e != CUDA_SUCCESS) return fl_fail(h, FLUID_E_CUDA, "cuStreamWaitValue32 failed (%d)", (int)e);
        }
        return 0;
    }
    if (!c->comm) return fl_fail(h, FLUID_E_COMM, "barrier: no communicator attached");
    FL_NCCL(h, c, c->AllReduce(c->d_token, c->d_token, 1, ncclFloat, ncclSum, c->comm, h->stream));
    h->launches++;
    return 0;
}

// ---- halo hand-over by peer stores ------------------------------------------------------------
// rank r's mapping of the buffer that corresponds to this rank's `local` (plane 0), or nullptr
static inline float* fl_peer_ptr(fluid_handle* h, int r, const float* local)
{
    SlabComm* c = (SlabComm*)h->comm;
    if (!c) return nullptr;
    for (int i = 0; i < FLUID_NFIELDS + 4; ++i)
        if (h->alloc_base[i] == local - FL_ZPAD * h->g.plane)
            return c->peer_alloc[r][i] ? c->peer_alloc[r][i] + FL_ZPAD * h->g.plane : nullptr;
    return nullptr;
}

bool fl_peer_mapped(fluid_handle* h, int r, const float* local) { return fl_peer_ptr(h, r, local) != nullptr; }

// collective, blocking: true on every rank iff `mine` is true on every rank (one-float ncclAllReduce; used once per
// handle for decisions that must not differ between ranks)
int fl_all_agree(fluid_handle* h, bool mine, bool* all)
{
    SlabComm* c = (SlabComm*)h->comm;
    if (!c || !c->comm) return fl_fail(h, FLUID_E_COMM, "agreement: no communicator attached");
    const float v = mine ? 1.0f : 0.0f;
    float sum = 0.0f;
    FL_CUDA(h, cudaMemcpyAsync(c->d_token, &v, sizeof v, cudaMemcpyHostToDevice, h->stream));
    FL_NCCL(h, c, c->AllReduce(c->d_token, c->d_token, 1, ncclFloat, ncclSum, c->comm, h->stream));
    FL_CUDA(h, cudaMemcpyAsync(&sum, c->d_token, sizeof sum, cudaMemcpyDeviceToHost, h->stream));
    FL_CUDA(h, cudaStreamSynchronize(h->stream));
    const float zero = 0.0f;                                              // the barrier's token starts from zero again
    FL_CUDA(h, cudaMemcpyAsync(c->d_token, &zero, sizeof zero, cudaMemcpyHostToDevice, h->stream));
    FL_CUDA(h, cudaStreamSynchronize(h->stream));
    *all = sum == (float)h->nranks;
    return 0;
}

// every buffer a sweep can write and the flag block are mapped from both z-neighbours, and the
// driver offers stream waits on memory: sweeps hand their halos over by peer stores
static inline bool fl_push_ready(fluid_handle* h);
bool fl_push_usable(fluid_handle* h) { return fl_push_ready(h); }
static inline bool fl_push_ready(fluid_handle* h)
{
    SlabComm* c = (SlabComm*)h->comm;
    if (!c || h->nranks == 1) return false;
    if (c->push_state == 0) {
        c->push_state = -1;
        const char* env = getenv("FLUID_SW_PUSH");
        bool ok = !(env && atoi(env) == 0);
        for (int side = 0; side < 2 && ok; ++side) {
            const int r = side == 0 ? h->rank - 1 : h->rank + 1;
            if (r < 0 || r >= h->nranks) continue;
            for (int i = 0; i < FLUID_IPC_COUNT && ok; ++i) ok = c->peer_alloc[r][i] != nullptr;
        }
        if (ok) {
            void* fn = nullptr;
            cudaDriverEntryPointQueryResult q;
            ok = cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &q) == cudaSuccess &&
                 q == cudaDriverEntryPointSuccess && fn;
            if (ok) {
                *(void**)(&c->WaitValue32) = fn;
                // probe: a wait that is already satisfied (spare word of the flag block, 0 >= 0)
                ok = c->WaitValue32((CUstream)h->stream, (CUdeviceptr)(h->d_sync + 8), 0u, CU_STREAM_WAIT_VALUE_GEQ) == CUDA_SUCCESS;
            }
        }
        if (ok) c->push_state = 1;
    }
    return c->push_state == 1;
}

struct PushArgs {
    const float4* src[8]; float4* dst[8];     // (field, side): nplanes consecutive planes each
    unsigned* flag[2];                        // the neighbours' arrival flags for this rank (nullptr = no neighbour)
    unsigned* counter;
    long long n4;                             // float4s per copy
    int ncopy; unsigned seq;
};

// Stores this rank's boundary planes into the neighbours' halo planes over NVLink and, once every
// CTA's stores are fenced, raises the neighbours' arrival flags to `seq`.
__global__ void __launch_bounds__(128) k_push_halo(const PushArgs a)
{
    for (int c = 0; c < a.ncopy; ++c) {
        const float4* __restrict__ s = a.src[c]; float4* __restrict__ d = a.dst[c];
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n4; i += (long long)gridDim.x * blockDim.x)
            d[i] = s[i];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned done = atomicAdd(a.counter, 1u);
        if (done == gridDim.x - 1) {                     // last CTA: all stores of all CTAs are fenced
            *a.counter = 0;
            __threadfence_system();
            for (int sd = 0; sd < 2; ++sd)
                if (a.flag[sd]) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.flag[sd]), "r"(a.seq) : "memory");
        }
    }
}

// hand the `nplanes` boundary planes per side of `fields` (the buffers a kernel just wrote) to the z-neighbours;
// seq = this hand-over's number in h->halo_seq order
int fl_halo_push(fluid_handle* h, int n, float* const* fields, cudaStream_t stream, unsigned seq, int nplanes)
{
    SlabComm* c = (SlabComm*)h->comm;
    const long long pl = h->g.plane;
    if (n > 4) return fl_fail(h, FLUID_E_ARG, "halo push: %d fields", n);
    PushArgs a = {};
    a.n4 = nplanes * pl / 4; a.seq = seq; a.counter = h->d_sync + 4;
    for (int side = 0; side < 2; ++side) {
        const int r = side == 0 ? h->rank - 1 : h->rank + 1;
        if (r < 0 || r >= h->nranks) continue;
        int zlo_r, dl_r; fl_slab_of(h->g.D, h->nranks, r, &zlo_r, &dl_r);
        // my planes 1 .. n -> lower neighbour's Dl_r+1 .. Dl_r+n; my Dl-n+1 .. Dl -> upper neighbour's 1-n .. 0.
        // The neighbour's flag [1] is "raised by its upper neighbour", [0] "by its lower neighbour".
        a.flag[side] = (unsigned*)c->peer_alloc[r][FLUID_IPC_SYNC] + (side == 0 ? 1 : 0);
        for (int i = 0; i < n; ++i) {
            float* peer = fl_peer_ptr(h, r, fields[i]);
            if (!peer) return fl_fail(h, FLUID_E_COMM, "halo push: buffer of rank %d is not IPC-mapped", r);
            a.src[a.ncopy] = (const float4*)(fields[i] + (side == 0 ? 1 : (long long)(h->g.Dl - nplanes + 1)) * pl);
            a.dst[a.ncopy] = (float4*)(peer + (side == 0 ? (long long)(dl_r + 1) : (long long)(1 - nplanes)) * pl);
            ++a.ncopy;
        }
    }
    const int ctas = nplanes > 2 ? 64 : 16;               // 128 threads, no shared memory: fits beside a resident sweep CTA
    k_push_halo<<<ctas, 128, 0, stream>>>(a);
    FL_CHECK_LAUNCH(h);
    return 0;
}

// where the fused sweep kernel writes: per face (0 lower, 1 upper) the neighbour's mapping of `fields`,
// its arrival flag for this rank and the neighbour-local index of the first halo plane
int fl_push_targets(fluid_handle* h, int n, float* const* fields, float* (*peer)[4], unsigned** flag, int* zdst)
{
    SlabComm* c = (SlabComm*)h->comm;
    for (int side = 0; side < 2; ++side) {
        flag[side] = nullptr; zdst[side] = 0;
        for (int i = 0; i < 4; ++i) peer[side][i] = nullptr;
        const int r = side == 0 ? h->rank - 1 : h->rank + 1;
        if (r < 0 || r >= h->nranks) continue;
        int zlo_r, dl_r; fl_slab_of(h->g.D, h->nranks, r, &zlo_r, &dl_r);
        zdst[side] = side == 0 ? dl_r + 1 : -1;
        flag[side] = (unsigned*)c->peer_alloc[r][FLUID_IPC_SYNC] + (side == 0 ? 1 : 0);
        for (int i = 0; i < n; ++i) {
            peer[side][i] = fl_peer_ptr(h, r, fields[i]);
            if (!peer[side][i]) return fl_fail(h, FLUID_E_COMM, "fused sweep: buffer of rank %d is not IPC-mapped", r);
        }
    }
    return 0;
}

// `stream` continues once both neighbours' halos of sweep `seq` have arrived
int fl_halo_wait(fluid_handle* h, cudaStream_t stream, unsigned seq)
{
    SlabComm* c = (SlabComm*)h->comm;
    for (int side = 0; side < 2; ++side) {
        const int r = side == 0 ? h->rank - 1 : h->rank + 1;
        if (r < 0 || r >= h->nranks) continue;
        CUresult e = c->WaitValue32((CUstream)stream, (CUdeviceptr)(h->d_sync + side), seq, CU_STREAM_WAIT_VALUE_GEQ);
        if (e != CUDA_SUCCESS) return fl_fail(h, FLUID_E_CUDA, "cuStreamWaitValue32 failed (%d)", (int)e);
    }
    return 0;
}

void fl_comm_release(fluid_handle* h)
{
    SlabComm* c = (SlabComm*)h->comm;
    if (!c) return;
    for (int r = 0; r < 8; ++r)
        for (int f = 0; f < FLUID_IPC_COUNT; ++f)
            if (c->peer_alloc[r][f]) cudaIpcCloseMemHandle(c->peer_alloc[r][f]);
    if (c->comm && c->CommDestroy) c->CommDestroy(c->comm);
    cudaFree(c->d_token);
    delete c;
    h->comm = nullptr;
}

// gather sources of advect(): every slab of prev_id, own pointer locally, IPC-mapped pointers for peers
int fl_gather_sources(fluid_handle* h, int prev_id, GatherSrc* out)
{
    out->n = h->nranks;
    if (h->nranks == 1) {
        out->base[0] = h->field[prev_id]; out->zlo[0] = h->g.zlo; out->dl[0] = h->g.Dl;
        return 0;
    }
    for (int r = 0; r < h->nranks; ++r) {
        fl_slab_of(h->g.D, h->nranks, r, &out->zlo[r], &out->dl[r]);
        if (r == h->rank) { out->base[r] = h->field[prev_id]; continue; }
        float* peer = fl_peer_ptr(h, r, h->field[prev_id]);
        if (!peer)
            return fl_fail(h, FLUID_E_COMM, "advect: field %d of rank %d is not IPC-mapped (fluid_ipc_import)", prev_id, r);
        out->base[r] = peer;
    }
    return 0;
}
