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
    float* peer_alloc[8][FLUID_NFIELDS] = {};              // IPC-mapped allocation bases (plane -1) of the other ranks
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

// Send my planes {1,2} down and {Dl-1,Dl} up, receive the neighbours' into {Dl+1,Dl+2} / {-1,0}.
int fl_halo_exchange_on(fluid_handle* h, int n, float* const* fields, cudaStream_t stream)
{
    if (h->nranks == 1) return 0;
    SlabComm* c = (SlabComm*)h->comm;
    if (!c || !c->comm) return fl_fail(h, FLUID_E_COMM, "halo exchange: no communicator attached to this slab handle (fluid_comm_init)");
    const size_t cnt = 2 * (size_t)h->g.plane;
    const long long pl = h->g.plane;
    const int Dl = h->g.Dl;
    FL_NCCL(h, c, c->GroupStart());
    for (int i = 0; i < n; ++i) {
        float* f = fields[i];
        if (h->rank > 0) {
            FL_NCCL(h, c, c->Send(f + 1 * pl, cnt, ncclFloat, h->rank - 1, c->comm, stream));
            FL_NCCL(h, c, c->Recv(f - 1 * pl, cnt, ncclFloat, h->rank - 1, c->comm, stream));
        }
        if (h->rank < h->nranks - 1) {
            FL_NCCL(h, c, c->Send(f + (long long)(Dl - 1) * pl, cnt, ncclFloat, h->rank + 1, c->comm, stream));
            FL_NCCL(h, c, c->Recv(f + (long long)(Dl + 1) * pl, cnt, ncclFloat, h->rank + 1, c->comm, stream));
        }
    }
    FL_NCCL(h, c, c->GroupEnd());
    h->launches++;
    return 0;
}

int fl_halo_exchange(fluid_handle* h, int n, float* const* fields) { return fl_halo_exchange_on(h, n, fields, h->stream); }

// all ranks' streams have reached this point before any continues (1-float all-reduce)
int fl_barrier(fluid_handle* h)
{
    if (h->nranks == 1) return 0;
    SlabComm* c = (SlabComm*)h->comm;
    if (!c || !c->comm) return fl_fail(h, FLUID_E_COMM, "barrier: no communicator attached");
    FL_NCCL(h, c, c->AllReduce(c->d_token, c->d_token, 1, ncclFloat, ncclSum, c->comm, h->stream));
    h->launches++;
    return 0;
}

void fl_comm_release(fluid_handle* h)
{
    SlabComm* c = (SlabComm*)h->comm;
    if (!c) return;
    for (int r = 0; r < 8; ++r)
        for (int f = 0; f < FLUID_NFIELDS; ++f)
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
    SlabComm* c = (SlabComm*)h->comm;
    for (int r = 0; r < h->nranks; ++r) {
        fl_slab_of(h->g.D, h->nranks, r, &out->zlo[r], &out->dl[r]);
        if (r == h->rank) { out->base[r] = h->field[prev_id]; continue; }
        if (!c || !c->peer_alloc[r][prev_id])
            return fl_fail(h, FLUID_E_COMM, "advect: field %d of rank %d is not IPC-mapped (fluid_ipc_import)", prev_id, r);
        out->base[r] = c->peer_alloc[r][prev_id] + h->g.plane;
    }
    return 0;
}
