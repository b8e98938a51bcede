// api_io.inl -- host <-> HBM transfers, sparse mutators, frame dumps, statistics.
// Included at the end of fluid_cuda.cu.

static int check_field(fluid_handle* h, int id, const char* who)
{
    if (!h) return FLUID_E_ARG;
    if (id < 0 || id >= FLUID_NFIELDS) return fl_fail(h, FLUID_E_ARG, "%s: bad field id %d", who, id);
    return 0;
}

// offset (floats) of local plane zl inside the GLOBAL reference volume
static size_t slab_host_offset(const fluid_handle* h, int zl) { return ref_plane_floats(h->g) * (size_t)(h->g.zlo - 1 + zl); }
// planes a slab holds data for (uploads fill all of them, halos included) ...
static int held_first(const fluid_handle* h) { return h->g.zl_lo; }
static int held_count(const fluid_handle* h) { return h->g.zl_hi - h->g.zl_lo + 1; }
// ... and the planes it OWNS (interior planes plus the global ghost plane(s) of the end slabs)
static int own_first(const fluid_handle* h) { return h->rank == 0 ? 0 : 1; }
static int own_count(const fluid_handle* h) { return (h->rank == h->nranks - 1 ? h->g.Dl + 1 : h->g.Dl) - own_first(h) + 1; }

extern "C" {

int fluid_upload_field(fluid_handle* h, int field_id, const float* padded_host)
{
    if (int e = check_field(h, field_id, "fluid_upload_field")) return e;
    if (!padded_host) return fl_fail(h, FLUID_E_ARG, "fluid_upload_field: null host pointer");
    const size_t pf = ref_plane_floats(h->g);
    const int zf = held_first(h), np = held_count(h);
    FL_CUDA(h, cudaMemcpyAsync(h->staging, padded_host + slab_host_offset(h, zf), pf * np * sizeof(float),
                               cudaMemcpyHostToDevice, h->stream));
    k_unpack<<<fl_grid_1d((long long)(pf * np)), 256, 0, h->stream>>>(h->g, zf, np, h->staging, h->field[field_id]);
    FL_CHECK_LAUNCH(h);
    if (field_id == FLUID_OBS) h->flags_dirty = true;
    // the staging buffer is reused by the next transfer: keep host-visible ordering simple
    FL_CUDA(h, cudaStreamSynchronize(h->stream));
    return 0;
}

int fluid_download_field(fluid_handle* h, int field_id, float* padded_host)
{
    if (int e = check_field(h, field_id, "fluid_download_field")) return e;
    if (!padded_host) return fl_fail(h, FLUID_E_ARG, "fluid_download_field: null host pointer");
    // a slab returns its OWNED planes only: interior planes plus the global ghost plane(s) it holds
    const size_t pf = ref_plane_floats(h->g);
    const int zf = own_first(h), np = own_count(h);
    k_pack<<<fl_grid_1d((long long)(pf * np)), 256, 0, h->stream>>>(h->g, zf, np, h->field[field_id], h->staging);
    FL_CHECK_LAUNCH(h);
    FL_CUDA(h, cudaMemcpyAsync(padded_host + slab_host_offset(h, zf), h->staging, pf * np * sizeof(float),
                               cudaMemcpyDeviceToHost, h->stream));
    FL_CUDA(h, cudaStreamSynchronize(h->stream));
    return 0;
}

int fluid_upload_obstacles(fluid_handle* h, const float* obs_padded)
{
    if (int e = fluid_upload_field(h, FLUID_OBS, obs_padded)) return e;
    return fl_rebuild_flags(h);
}

static int sparse_stage(fluid_handle* h, int n, const int* xyz, const float* val, int nval)
{
    if (n > h->sparse_cap) {
        cudaFree(h->d_xyz); cudaFree(h->d_val); h->d_xyz = nullptr; h->d_val = nullptr;
        int cap = n < 4096 ? 4096 : n;
        FL_CUDA(h, cudaMalloc(&h->d_xyz, (size_t)cap * 3 * sizeof(int)));
        FL_CUDA(h, cudaMalloc(&h->d_val, (size_t)cap * 3 * sizeof(float)));
        h->sparse_cap = cap;
    }
    FL_CUDA(h, cudaMemcpyAsync(h->d_xyz, xyz, (size_t)n * 3 * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    if (val) FL_CUDA(h, cudaMemcpyAsync(h->d_val, val, (size_t)n * nval * sizeof(float),
                                        cudaMemcpyHostToDevice, h->stream));
    return 0;
}

static int check_xyz(fluid_handle* h, int n, const int* xyz, const char* who)
{
    if (!h) return FLUID_E_ARG;
    if (n < 0 || (n > 0 && !xyz)) return fl_fail(h, FLUID_E_ARG, "%s: bad arguments", who);
    for (int i = 0; i < n; ++i)
        if (xyz[3 * i] < 0 || xyz[3 * i] > h->g.W + 1 || xyz[3 * i + 1] < 0 || xyz[3 * i + 1] > h->g.H + 1 ||
            xyz[3 * i + 2] < 0 || xyz[3 * i + 2] > h->g.D + 1)
            return fl_fail(h, FLUID_E_ARG, "%s: cell %d (%d,%d,%d) outside the padded grid", who, i,
                           xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    return 0;
}

int fluid_add_obstacles(fluid_handle* h, int n, const int* xyz)
{
    if (int e = check_xyz(h, n, xyz, "fluid_add_obstacles")) return e;
    if (n == 0) return 0;
    if (int e = sparse_stage(h, n, xyz, nullptr, 0)) return e;
    k_add_obstacles<<<(n + 255) / 256, 256, 0, h->stream>>>(h->g, h->field[FLUID_OBS], n, h->d_xyz);
    FL_CHECK_LAUNCH(h);
    h->flags_dirty = true;
    return 0;
}

int fluid_voxelize(fluid_handle* h, const float* tris, int n_tris, float scale, float rot_x, float rot_y,
                   float rot_z, float tx, float ty, float tz, int sampling, long long* added)
{
    if (!h) return FLUID_E_ARG;
    return fl_voxelize(h, tris, n_tris, scale, rot_x, rot_y, rot_z, tx, ty, tz, sampling, added);
}

int fluid_render_slice(fluid_handle* h, int field_id, int z, double vmin, double vmax, double obstacle_alpha,
                       unsigned char* rgb_host)
{
    return fl_render_slice(h, field_id, z, vmin, vmax, obstacle_alpha, rgb_host);
}

int fluid_density_cmap(double lut[768])
{
    if (!lut) return FLUID_E_ARG;
    pp_density_cmap(reinterpret_cast<double(*)[3]>(lut));
    return 0;
}

int fluid_obstacle_mesh(fluid_handle* h, double level, long long max_verts, long long max_faces, long long* n_verts,
                        long long* n_faces, float* verts, int* faces)
{
    return fl_obstacle_mesh(h, level, max_verts, max_faces, n_verts, n_faces, verts, faces);
}

int fluid_streamlines(fluid_handle* h, const fluid_streamline_params* params, int max_lines, long long max_points,
                      int* n_lines, long long* n_points, int* line_len, double* norm_speed, double* points)
{
    return fl_streamlines(h, params, max_lines, max_points, n_lines, n_points, line_len, norm_speed, points);
}

// Batched mutators keep the reference's per-call semantics when a cell appears more than once in a batch
// (simulation.cpp:155-178 runs the calls one after the other): occurrence[i] = how many earlier entries name the
// same cell.  A batch in strictly increasing cell order (the usual loop nest) has none and skips the hashing.
static int occurrences(const fluid_handle* h, int n, const int* xyz, std::vector<int>* occ)
{
    const long long Wp = h->g.W + 2, Hp = h->g.H + 2;
    auto lin = [&](int i) { return xyz[3 * i] + Wp * (xyz[3 * i + 1] + Hp * (long long)xyz[3 * i + 2]); };
    bool increasing = true;
    for (int i = 1; i < n && increasing; ++i) increasing = lin(i) > lin(i - 1);
    if (increasing) return 0;
    std::unordered_map<long long, int> seen;
    seen.reserve((size_t)n * 2);
    occ->assign((size_t)n, 0);
    int most = 0;
    for (int i = 0; i < n; ++i) { const int k = seen[lin(i)]++; (*occ)[i] = k; if (k > most) most = k; }
    return most;
}

int fluid_add_density(fluid_handle* h, int n, const int* xyz, const float* amount)
{
    if (int e = check_xyz(h, n, xyz, "fluid_add_density")) return e;
    if (n == 0) return 0;
    if (!amount) return fl_fail(h, FLUID_E_ARG, "fluid_add_density: null amounts");
    std::vector<int> occ;
    const int most = occurrences(h, n, xyz, &occ);
    if (most == 0) {
        if (int e = sparse_stage(h, n, xyz, amount, 1)) return e;
        k_add_density<<<(n + 255) / 256, 256, 0, h->stream>>>(h->g, h->field[FLUID_DENS], n, h->d_xyz, h->d_val);
        FL_CHECK_LAUNCH(h);
        return 0;
    }
    // repeated cells: round k adds the k-th mention of every cell, so each cell's float additions happen in call
    // order (one addition per cell and launch: the atomicAdd is then an ordinary, deterministic addition)
    std::vector<int> xs; std::vector<float> am;
    for (int k = 0; k <= most; ++k) {
        xs.clear(); am.clear();
        for (int i = 0; i < n; ++i) if (occ[i] == k) { xs.insert(xs.end(), xyz + 3 * i, xyz + 3 * i + 3); am.push_back(amount[i]); }
        const int m = (int)am.size();
        if (int e = sparse_stage(h, m, xs.data(), am.data(), 1)) return e;
        k_add_density<<<(m + 255) / 256, 256, 0, h->stream>>>(h->g, h->field[FLUID_DENS], m, h->d_xyz, h->d_val);
        FL_CHECK_LAUNCH(h);
        FL_CUDA(h, cudaStreamSynchronize(h->stream));       // the host staging vectors are reused by the next round
    }
    return 0;
}

int fluid_set_velocity(fluid_handle* h, int n, const int* xyz, const float* v)
{
    if (int e = check_xyz(h, n, xyz, "fluid_set_velocity")) return e;
    if (n == 0) return 0;
    if (!v) return fl_fail(h, FLUID_E_ARG, "fluid_set_velocity: null values");
    std::vector<int> occ;
    if (occurrences(h, n, xyz, &occ) > 0) {
        // repeated cells: the last call wins (simulation.cpp:172-177), so only each cell's last mention is staged
        std::vector<int> last((size_t)n, 0), xs; std::vector<float> vs;
        {
            const long long Wp = h->g.W + 2, Hp = h->g.H + 2;
            std::unordered_map<long long, int> at;
            for (int i = 0; i < n; ++i) at[xyz[3 * i] + Wp * (xyz[3 * i + 1] + Hp * (long long)xyz[3 * i + 2])] = i;
            for (auto& kv : at) last[(size_t)kv.second] = 1;
        }
        for (int i = 0; i < n; ++i) if (last[i]) { xs.insert(xs.end(), xyz + 3 * i, xyz + 3 * i + 3); vs.insert(vs.end(), v + 3 * i, v + 3 * i + 3); }
        const int m = (int)xs.size() / 3;
        if (int e = sparse_stage(h, m, xs.data(), vs.data(), 3)) return e;
        k_set_velocity<<<(m + 255) / 256, 256, 0, h->stream>>>(h->g, h->field[FLUID_VX], h->field[FLUID_VY],
                                                              h->field[FLUID_VZ], m, h->d_xyz, h->d_val);
        FL_CHECK_LAUNCH(h);
        FL_CUDA(h, cudaStreamSynchronize(h->stream));       // the host staging vectors go out of scope
        return 0;
    }
    if (int e = sparse_stage(h, n, xyz, v, 3)) return e;
    k_set_velocity<<<(n + 255) / 256, 256, 0, h->stream>>>(h->g, h->field[FLUID_VX], h->field[FLUID_VY],
                                                          h->field[FLUID_VZ], n, h->d_xyz, h->d_val);
    FL_CHECK_LAUNCH(h);
    return 0;
}

// frame write of step(), simulation.cpp:140-148: dens, obs, v_x, v_y, v_z.  One kernel on the compute
// stream takes the snapshot (all requested fields repacked to the reference layout, HBM to HBM), then one
// D2H copy per field runs on the dump stream so the next time step overlaps the PCIe transfer.
int fluid_dump_frame_async(fluid_handle* h, float* const dst[5])
{
    if (!h || !dst) return FLUID_E_ARG;
    static const int ids[5] = { FLUID_DENS, FLUID_OBS, FLUID_VX, FLUID_VY, FLUID_VZ };
    const size_t pf = ref_plane_floats(h->g);
    const int zf = own_first(h), np = own_count(h);
    const size_t nfl = pf * np;
    for (int i = 0; i < 5; ++i)
        if (dst[i] && !h->frame_staging[i]) FL_CUDA(h, cudaMalloc(&h->frame_staging[i], nfl * sizeof(float)));
    // the previous dump must have left the staging buffers before they are overwritten
    if (h->frame_in_flight) FL_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_frame_done, 0));
    PackBatch pb = {}; int nf = 0;
    for (int i = 0; i < 5; ++i) {
        if (!dst[i]) continue;
        pb.src[nf] = h->field[ids[i]]; pb.dst[nf] = h->frame_staging[i]; ++nf;
    }
    if (nf > 0) {                                         // one snapshot launch for the whole frame
        const dim3 blk(32, 8), grd((unsigned)((h->g.rows * np + 7) / 8), (unsigned)nf);
        k_pack_rows<<<grd, blk, 0, h->stream>>>(h->g, zf, np, pb);
        FL_CHECK_LAUNCH(h);
    }
    FL_CUDA(h, cudaEventRecord(h->ev_frame_packed, h->stream));
    FL_CUDA(h, cudaStreamWaitEvent(h->dump_stream, h->ev_frame_packed, 0));
    for (int i = 0; i < 5; ++i) {
        if (!dst[i]) continue;
        FL_CUDA(h, cudaMemcpyAsync(dst[i] + slab_host_offset(h, zf), h->frame_staging[i], nfl * sizeof(float),
                                   cudaMemcpyDeviceToHost, h->dump_stream));
    }
    FL_CUDA(h, cudaEventRecord(h->ev_frame_done, h->dump_stream));
    h->frame_in_flight = true;
    return 0;
}

int fluid_frame_wait(fluid_handle* h)
{
    if (!h) return FLUID_E_ARG;
    if (h->frame_in_flight) { FL_CUDA(h, cudaEventSynchronize(h->ev_frame_done)); h->frame_in_flight = false; }
    return 0;
}

// statistics of run() (simulation.cpp:73-90) over the padded cells this slab owns
int fluid_stats(fluid_handle* h, double out[9])
{
    if (!h || !out) return FLUID_E_ARG;
    const int first = (h->rank == 0) ? 0 : 1;
    const int last = (h->rank == h->nranks - 1) ? h->g.Dl + 1 : h->g.Dl;
    const long long nq = (h->g.plane >> 2) * (last - first + 1);
    const int nb = fl_grid_1d(nq, 256, h->n_partial);
    k_stats_quads<<<nb, 256, 0, h->stream>>>(h->g, first, last, h->field[FLUID_DENS], h->field[FLUID_VX],
                                             h->field[FLUID_VY], h->field[FLUID_VZ], h->d_partial,
                                             reinterpret_cast<unsigned*>(h->d_out + FL_NSTAT), h->d_out);
    FL_CHECK_LAUNCH(h);
    FL_CUDA(h, cudaMemcpyAsync(out, h->d_out, FL_NSTAT * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    FL_CUDA(h, cudaStreamSynchronize(h->stream));
    return 0;
}

int fluid_divergence_norm(fluid_handle* h, double out[2])
{
    if (!h || !out) return FLUID_E_ARG;
    if (int e = fl_rebuild_flags(h)) return e;
    const long long n = (long long)h->g.W * h->g.H * h->g.Dl;
    const int nb = fl_grid_1d(n, 256, h->n_partial);
    k_divnorm_partial<<<nb, 256, 0, h->stream>>>(h->g, h->field[FLUID_DIVERGENCE], h->cflags, h->d_partial);
    FL_CHECK_LAUNCH(h);
    std::vector<double> part(2 * (size_t)nb);
    FL_CUDA(h, cudaMemcpyAsync(part.data(), h->d_partial, part.size() * sizeof(double),
                               cudaMemcpyDeviceToHost, h->stream));
    FL_CUDA(h, cudaStreamSynchronize(h->stream));
    double s = 0.0, m = 0.0;
    for (int i = 0; i < nb; ++i) { s += part[2 * i]; if (part[2 * i + 1] > m) m = part[2 * i + 1]; }
    out[0] = sqrt(s); out[1] = m;      // slab mode: caller combines (sum of squares is out[0]^2)
    return 0;
}

}  // extern "C"
