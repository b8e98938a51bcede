// api_ops.inl -- fluid_step / fluid_call and the slab-mode hooks.  Included by fluid_cuda.cu.

// ---- slab-mode hooks (multi-GPU z-slabs) ---------------------------------------
// Single-GPU handles never reach these with nranks > 1.
int fl_halo_exchange(fluid_handle* h, int n, float* const* fields)
{
    (void)n; (void)fields;
    if (h->nranks == 1) return 0;
    return fl_fail(h, FLUID_E_COMM, "halo exchange: no communicator attached to this slab handle");
}

void fl_comm_release(fluid_handle* h) { (void)h; }

int fl_gather_sources(fluid_handle* h, int prev_id, GatherSrc* out)
{
    if (h->nranks == 1) {
        out->n = 1; out->base[0] = h->field[prev_id]; out->zlo[0] = h->g.zlo; out->dl[0] = h->g.Dl;
        return 0;
    }
    return fl_fail(h, FLUID_E_COMM, "advect: slab gather sources not attached");
}

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
        case FLUID_OP_SET_BOUNDS:  return fl_set_bounds(h, b, field_id);
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
    }
    return fl_fail(h, FLUID_E_ARG, "fluid_call: unknown op %d", op);
}

}  // extern "C"
