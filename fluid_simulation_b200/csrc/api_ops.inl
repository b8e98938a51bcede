// api_ops.inl -- fluid_step / fluid_call and the slab-mode hooks.  Included by fluid_cuda.cu.

// ---- slab-mode entry points (multi-GPU z-slabs); the mechanisms live in comm.cuh -------
extern "C" {

// n x { inlet density, buffer = dens, step() } -- the body of run()'s loop (simulation.cpp:63-71)
int fluid_step(fluid_handle* h, int n_steps)
{
    if (!h || n_steps < 0) return fl_fail(h, FLUID_E_ARG, "fluid_step: bad arguments");
    for (int i = 0; i < n_steps; ++i)
        if (int e = fl_run_iteration(h)) return e;
    return 0;
}

int fluid_call(fluid_handle* h, int op, int b, int field_id, int prev_id, float a, float c)
{
    if (!h) return FLUID_E_ARG;
    const bool needs_field = op == FLUID_OP_SET_BOUNDS || op == FLUID_OP_LIN_SOLVE ||
                             op == FLUID_OP_DIFFUSE || op == FLUID_OP_ADVECT || op == FLUID_OP_COPY;
    const bool needs_prev = op == FLUID_OP_LIN_SOLVE || op == FLUID_OP_DIFFUSE || op == FLUID_OP_ADVECT || op == FLUID_OP_COPY;
    if (needs_field && (field_id < 0 || field_id >= FLUID_NFIELDS || b < 0 || b > 3))
        return fl_fail(h, FLUID_E_ARG, "fluid_call(op=%d): bad field %d or b %d", op, field_id, b);
    if (needs_prev && (prev_id < 0 || prev_id >= FLUID_NFIELDS || prev_id == field_id))
        return fl_fail(h, FLUID_E_ARG, "fluid_call(op=%d): bad prev field %d", op, prev_id);
    switch (op) {
        case FLUID_OP_SET_BOUNDS: {
            if (int e = fl_set_bounds(h, b, field_id)) return e;
            float* f[1] = { h->field[field_id] };
            return fl_halo_exchange(h, 1, f);
        }
        case FLUID_OP_LIN_SOLVE:   return fl_lin_solve(h, 1, &b, &field_id, &prev_id, a, c);
        case FLUID_OP_DIFFUSE:     return fl_diffuse(h, 1, &b, &field_id, &prev_id);
        case FLUID_OP_DIVERGENCE:  return fl_divergence(h);
        case FLUID_OP_GRADIENT:    return fl_gradient(h);
        case FLUID_OP_PROJECT:     return fl_project(h);
        case FLUID_OP_ADVECT:      return fl_advect(h, b, field_id, prev_id);
        case FLUID_OP_INFLOW_DENS: return fl_inflow_density(h);
        case FLUID_OP_INFLOW_VEL:  return fl_inflow_velocity(h);
        case FLUID_OP_STEP:        return fl_step_only(h);
        case FLUID_OP_COPY:        return fl_copy_field(h, field_id, prev_id);
        case FLUID_OP_ADVECT_VEL:  return fl_advect_velocity(h);
    }
    return fl_fail(h, FLUID_E_ARG, "fluid_call: unknown op %d", op);
}


// ncclGetUniqueId for rank 0; the host program broadcasts the 128 bytes to the other ranks
int fluid_comm_unique_id(const char* nccl_lib_path, char out[128])
{
    static SlabComm loader;
    std::string err;
    if (comm_load(&loader, nccl_lib_path, &err)) return fl_fail(nullptr, FLUID_E_COMM, "%s", err.c_str());
    ncclUniqueId id;
    ncclResult_t r = loader.GetUniqueId(&id);
    if (r != ncclSuccess) return fl_fail(nullptr, FLUID_E_COMM, "ncclGetUniqueId: %s", loader.GetErrorString(r));
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    memcpy(out, &id, 128);
    return 0;
}

int fluid_comm_init(fluid_handle* h, const char* nccl_lib_path, const char id_bytes[128])
{
    if (!h || !id_bytes) return FLUID_E_ARG;
    if (h->nranks == 1) return 0;
    if (!h->comm) h->comm = new SlabComm();
    SlabComm* c = (SlabComm*)h->comm;
    std::string err;
    if (comm_load(c, nccl_lib_path, &err)) return fl_fail(h, FLUID_E_COMM, "%s", err.c_str());
    ncclUniqueId id; memcpy(&id, id_bytes, 128);
    FL_NCCL(h, c, c->CommInitRank(&c->comm, h->nranks, id, h->rank));
    FL_CUDA(h, cudaMalloc(&c->d_token, sizeof(float)));
    FL_CUDA(h, cudaMemsetAsync(c->d_token, 0, sizeof(float), h->stream));
    return 0;
}

int fluid_ipc_export(fluid_handle* h, int alloc_id, char out[64])
{
    if (!h || !out) return FLUID_E_ARG;
    if (alloc_id < 0 || alloc_id >= FLUID_IPC_COUNT)
        return fl_fail(h, FLUID_E_ARG, "fluid_ipc_export: bad allocation id %d", alloc_id);
    // a field id names the buffer the field lives in right now (the solver permutes them)
    void* base = alloc_id == FLUID_IPC_SYNC ? (void*)h->d_sync
               : alloc_id >= FLUID_IPC_SCRATCH0 ? (void*)(h->scratch[alloc_id - FLUID_IPC_SCRATCH0] - FL_ZPAD * h->g.plane)
               : (void*)(h->field[alloc_id] - FL_ZPAD * h->g.plane);
    cudaIpcMemHandle_t mh;
    FL_CUDA(h, cudaIpcGetMemHandle(&mh, base));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t size");
    memcpy(out, &mh, 64);
    return 0;
}

int fluid_ipc_import(fluid_handle* h, int peer_rank, int alloc_id, const char in[64])
{
    if (!h || !in) return FLUID_E_ARG;
    if (alloc_id < 0 || alloc_id >= FLUID_IPC_COUNT)
        return fl_fail(h, FLUID_E_ARG, "fluid_ipc_import: bad allocation id %d", alloc_id);
    if (peer_rank < 0 || peer_rank >= h->nranks || peer_rank == h->rank)
        return fl_fail(h, FLUID_E_ARG, "fluid_ipc_import: bad peer rank %d", peer_rank);
    if (!h->comm) h->comm = new SlabComm();
    SlabComm* c = (SlabComm*)h->comm;
    cudaIpcMemHandle_t mh; memcpy(&mh, in, 64);
    void* p = nullptr;
    FL_CUDA(h, cudaIpcOpenMemHandle(&p, mh, cudaIpcMemLazyEnablePeerAccess));
    // Imports happen at attach time, before any solve has permuted the buffers, and every rank
    // permutes alike: the peer's buffer with id i keeps corresponding to this rank's alloc_base[i].
    int slot = alloc_id;
    if (alloc_id < FLUID_NFIELDS)
        for (int i = 0; i < FLUID_NFIELDS + 4; ++i)
            if (h->alloc_base[i] == h->field[alloc_id] - FL_ZPAD * h->g.plane) slot = i;
    if (alloc_id >= FLUID_IPC_SCRATCH0 && alloc_id < FLUID_IPC_SYNC)
        for (int i = 0; i < FLUID_NFIELDS + 4; ++i)
            if (h->alloc_base[i] == h->scratch[alloc_id - FLUID_IPC_SCRATCH0] - FL_ZPAD * h->g.plane) slot = i;
    c->peer_alloc[peer_rank][slot] = (float*)p;
    return 0;
}

int fluid_barrier(fluid_handle* h)
{
    if (!h) return FLUID_E_ARG;
    return fl_barrier(h);
}

}  // extern "C"
