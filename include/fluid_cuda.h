/* include/fluid_cuda.h -- C ABI of libfluid_cuda, the B200 (sm_100a) drop-in for the
 * per-timestep hot path of the reference wind-tunnel solver.
 *
 * The reference has no FFI layer: its boundary is the C++ class `Simulation`
 * (/root/reference/simulation.h:11-91).  A host `Simulation` with the same public
 * API (fluid_simulation_b200/host/simulation.h) owns a `fluid_handle*` instead of
 * eleven std::vector<float> and forwards to the entry points below; each entry
 * point cites the reference code it replaces.
 *
 * Conventions
 *  - plain C types only; every call returns 0 on success or a negative FLUID_E_*
 *    code and never throws; `fluid_last_error` gives the text.
 *  - host arrays use the REFERENCE layout: (w+2)*(h+2)*(d+2) floats, index
 *    x + y*(w+2) + z*(w+2)*(h+2), x fastest (simulation.h:9 `global_at`),
 *    coordinates 1..dim interior, 0 and dim+1 ghosts.
 *  - the handle owns all device memory and streams; fields stay resident in HBM.
 *  - there is NO CPU fallback: every entry point fails with FLUID_E_CUDA when no
 *    sm_100-class device is usable.
 *  - one host thread drives a handle; calls are asynchronous on the handle's
 *    stream unless stated, `fluid_sync` / downloads / stats block.
 */
#ifndef FLUID_CUDA_H
#define FLUID_CUDA_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fluid_handle fluid_handle;

/* field ids = the 11 state vectors of simulation.h:16-27 */
enum fluid_field {
    FLUID_VX = 0, FLUID_VY = 1, FLUID_VZ = 2,
    FLUID_VX_PREV = 3, FLUID_VY_PREV = 4, FLUID_VZ_PREV = 5,
    FLUID_DENS = 6, FLUID_BUFFER = 7, FLUID_OBS = 8,
    FLUID_PRESSURE = 9, FLUID_DIVERGENCE = 10, FLUID_NFIELDS = 11
};

/* creation flags */
enum fluid_flags {
    FLUID_SOLVER_RB     = 0x0,  /* default: red-black Gauss-Seidel, colour (i+j+l)&1==0 first        */
    FLUID_SOLVER_JACOBI = 0x1,  /* Jacobi sweeps                                                      */
    FLUID_SOLVER_LEX    = 0x2,  /* the reference's lexicographic in-place order (simulation.cpp:260),
                                   run as a hyperplane wavefront: bit-identical to the 1-thread CPU   */
    FLUID_SOLVER_MASK   = 0x3,
    FLUID_ELIDE_DEAD    = 0x10, /* skip diffuse(0,dens,buffer) (simulation.cpp:135): its result is
                                   fully overwritten by advect (:136); output is bit-identical         */
    FLUID_NO_TILED      = 0x20, /* force the simple global-memory sweep kernels (debug / A-B checks)  */
    FLUID_TILED_ALWAYS  = 0x40  /* use the TMA/shared-memory sweep kernel even on tiny grids (tests)  */
};

/* per-function parity hooks for fluid_call (the reference's private methods, simulation.h:33-40) */
enum fluid_op {
    FLUID_OP_SET_BOUNDS = 1,   /* setBounds(b, field)                       simulation.cpp:183 */
    FLUID_OP_LIN_SOLVE  = 2,   /* linearSolver(b, field, prev, a, c)        simulation.cpp:251 */
    FLUID_OP_DIFFUSE    = 3,   /* diffuse(b, field, prev)                   simulation.cpp:278 */
    FLUID_OP_DIVERGENCE = 4,   /* project() part 1: div, p=0, 2x setBounds  simulation.cpp:295-319 */
    FLUID_OP_GRADIENT   = 5,   /* project() part 3: v -= grad p, 3x bounds  simulation.cpp:322-361 */
    FLUID_OP_PROJECT    = 6,   /* project(v_x,v_y,v_z,pressure,divergence)  simulation.cpp:289 */
    FLUID_OP_ADVECT     = 7,   /* advect(b, field, prev)                    simulation.cpp:367 */
    FLUID_OP_INFLOW_DENS = 8,  /* addDensity(1,j,k,0.001f) on the x=1 face  simulation.cpp:65-67 */
    FLUID_OP_INFLOW_VEL  = 9,  /* setVelocity(1,j,k,speed,0,0) on x=1 face  simulation.cpp:103-105 */
    FLUID_OP_STEP        = 10, /* Simulation::step() without run()'s prologue simulation.cpp:96-138 */
    FLUID_OP_COPY        = 11, /* field = prev (buffer = dens, v_prev = v)    simulation.cpp:70, 108-110 */
    FLUID_OP_ADVECT_VEL  = 12  /* advect(1,v_x,..); advect(2,v_y,..); advect(3,v_z,..) as step() runs them  simulation.cpp:125-127 */
};

enum fluid_error {
    FLUID_OK = 0, FLUID_E_ARG = -1, FLUID_E_CUDA = -2, FLUID_E_NOMEM = -3,
    FLUID_E_STATE = -4, FLUID_E_COMM = -5
};

/* Simulation::Simulation (simulation.cpp:17-44): allocates the 11 zeroed fields in HBM on the
 * CURRENT CUDA device.  Returns NULL on failure (text via fluid_last_error(NULL)). */
fluid_handle* fluid_create(int w, int h, int d, int speed, float dt, float diff, float visc,
                           int acc, unsigned flags);

/* z-slab member of a multi-GPU run: this process owns global planes z_lo..z_lo+d_local-1 of a
 * w x h x d tunnel (the reference has no distributed mode).  `rank`/`nranks` order the
 * slabs along z (equal slabs, remainder planes to the low ranks; d >= 4*nranks: a slab keeps four halo
 * planes per interior side and every slab must hold at least as many).  Before the first
 * device operation the host program attaches the communicator (fluid_comm_init) and maps the peers'
 * advection sources (fluid_ipc_export / fluid_ipc_import); see fluid_simulation_b200/multi_gpu.py. */
fluid_handle* fluid_create_slab(int w, int h, int d, int speed, float dt, float diff, float visc,
                                int acc, unsigned flags, int rank, int nranks);

void fluid_destroy(fluid_handle*);

/* parameters that the reference exposes as public data members (simulation.h:44-54) */
int fluid_set_acc(fluid_handle*, int acc);
/* speed, dt, diff, visc, acc are public and mutable in the reference (simulation.h:44-54: `sim.dt = ...` takes effect
 * at the next step()); a host class that keeps them as members pushes them here before every step / run */
int fluid_set_params(fluid_handle*, int speed, float dt, float diff, float visc, int acc);
int fluid_get_dims(fluid_handle*, int* w, int* h, int* d, int* z_lo, int* d_local);

/* whole-field transfers in reference layout (replace direct access to the std::vectors).
 * In slab mode the host array is still the GLOBAL padded volume; each rank moves its planes. */
int fluid_upload_field(fluid_handle*, int field_id, const float* padded_host);
int fluid_download_field(fluid_handle*, int field_id, float* padded_host);        /* blocking */

/* addObstacle (simulation.cpp:155) in bulk: the padded obs volume (0.0f / 1.0f), uploaded once;
 * also builds the per-cell flag volume used by every kernel. */
int fluid_upload_obstacles(fluid_handle*, const float* obs_padded);

/* sparse mutators, flushed in order before the next device operation:
 * addObstacle / addDensity / setVelocity (simulation.cpp:155-178), n cells at once. */
int fluid_add_obstacles(fluid_handle*, int n, const int* xyz /* 3n */);
int fluid_add_density(fluid_handle*, int n, const int* xyz /* 3n */, const float* amount /* n */);
int fluid_set_velocity(fluid_handle*, int n, const int* xyz /* 3n */, const float* v /* 3n */);

/* loadSTLIntoObstacles (object_loader.cpp:288-449) from the triangle list on: rotate the mesh about the
 * reference's objCenter, build the 64^3 rejection grid, ray-parity test every lattice sample against every
 * triangle and set obs = 1 in the simulation cells the inside samples map to; `added` receives the
 * reference's "Added N obstacle points" count.  `tris` = n_tris x 9 host floats (v1, v2, v3 as read from
 * the STL, object_loader.cpp:98-174).  The reference seeds its per-sample jitter and ray direction from
 * the OpenMP thread id (:399), so its mask is not reproducible; `sampling` picks the replacement:
 * HASHED = same distributions from a hash of the sample index, FIXED = no jitter and ray (1, 0.1, 0.2),
 * CELLS = one sample per simulation cell instead of the 200^3 lattice (no porous masks).  Blocking. */
enum { FLUID_VOX_HASHED = 0, FLUID_VOX_FIXED = 1, FLUID_VOX_CELLS = 2 };
int fluid_voxelize(fluid_handle*, const float* tris, int n_tris, float scale, float rot_x, float rot_y,
                   float rot_z, float translate_x, float translate_y, float translate_z, int sampling,
                   long long* added);

/* n x { inlet density, buffer = dens, step() } = the body of run()'s loop (simulation.cpp:65-71),
 * without the frame write.  Asynchronous. */
int fluid_step(fluid_handle*, int n_steps);

/* one private method of the reference on resident fields (parity hooks). a/c only for LIN_SOLVE. */
int fluid_call(fluid_handle*, int op, int b, int field_id, int prev_id, float a, float c);

/* frame write of step() (simulation.cpp:140-148): dens, obs, v_x, v_y, v_z in reference layout,
 * copied on a side stream into `dst[i]` (pinned host memory for overlap; NULL entries skipped).
 * Ordered after all work submitted so far; returns immediately.  fluid_frame_wait blocks until
 * the last dump has landed. */
int fluid_dump_frame_async(fluid_handle*, float* const dst[5]);
int fluid_frame_wait(fluid_handle*);

/* run()'s statistics (simulation.cpp:73-90) over all padded cells:
 * out = { sum(dens), min(dens), max(dens), min(v_x), max(v_x), min(v_y), max(v_y), min(v_z), max(v_z) } */
int fluid_stats(fluid_handle*, double out[9]);

/* sqrt(sum div^2) and max|div| of the stored divergence field over interior fluid cells */
int fluid_divergence_norm(fluid_handle*, double out[2]);

/* fluid_stats without a handle and WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  dens, v_x, v_y, v_z: DEVICE pointers to n_cells floats each
 * (e.g. the reference's padded arrays, (w+2)(h+2)(d+2) cells); out_device: DEVICE pointer to 9 doubles, laid out
 * as fluid_stats' out; stream: a cudaStream_t of the current device (NULL = the default stream).  Asynchronous:
 * two kernels are queued on `stream`, nothing is synchronised, and no cudaMalloc / cudaFree or other allocating
 * routine is called.  The per-block partials live in a file-scope device array of 4096 x 9 doubles (288 KiB
 * that every GPU loading the library holds until the process exits), one block per 65536 cells, hence
 *     n_cells <= FLUID_STATS_STATIC_MAX_CELLS = 2^28;  larger (or n_cells < 1) fails with FLUID_E_ARG
 * and a text in fluid_last_error(NULL).  The array is SHARED by all calls on a device: calls that overlap on
 * the GPU (different streams, different threads) race on it -- order them on one stream or with events.
 * Nothing in it has to start at zero (each call overwrites every partial it reads).  The first call on a
 * device looks the array's device address up (cudaGetSymbolAddress): make that call before a capture begins. */
#define FLUID_STATS_STATIC_MAX_CELLS (1LL << 28)
int fluid_stats_static(const float* dens, const float* v_x, const float* v_y, const float* v_z,
                       long long n_cells, double* out_device, void* stream);

/* ---- the offline viewers' arithmetic on the resident fields; single-GPU handles only.
 * The viewers' Qt windows and OpenGL draw calls are not reproduced: these return what they would draw. ---- */

/* gui.py:61-79 + 265-300 (FluidViewer.update_image): plane z (0..d+1) of `field_id` colour-mapped with the
 * viewers' density_cmap (Normalize(vmin, vmax, clip=True), 256 entries) to RGB888 and darkened by
 * (1 - obstacle_alpha) where obs > 0.5 (the viewer uses 0.2).  rgb_host receives (h+2) rows x (w+2) columns x 3
 * bytes, row y, column x, exactly the array the viewer hands to QImage.  Blocking. */
int fluid_render_slice(fluid_handle*, int field_id, int z, double vmin, double vmax, double obstacle_alpha,
                       unsigned char* rgb_host);

/* the 256 x 3 float64 table behind that colour map (LinearSegmentedColormap.from_list over the seven named
 * colours of gui.py:38-41 / GUI/config.py:14-17); host-only, needs no device */
int fluid_density_cmap(double lut[768]);

/* GUI/utils.py:118-214 generate_streamlines on the current v_x, v_y, v_z, obs (float64 arithmetic on the
 * float32 fields, padded [x, y, z] coordinates as in GUI/main_window.py:227-230): seeds on the
 * density x density/2 x density/2 lattice near the obstacle's bounding box, unit-speed integration of
 * integration_steps/2 steps backward and forward (:83-115), the length / velocity-change / proximity
 * filters (:173-198).  Accepted lines come back in the reference's order: line_len[i] points each, packed
 * x, y, z in `points`; norm_speed[i] = min(max |v| on the line / (max(v_x, v_y, v_z) + 1e-6), 1), the
 * argument of the viewers' colour map (:203-208).  n_lines / n_points always receive the totals; if they
 * exceed max_lines / max_points the call fails with FLUID_E_ARG and copies nothing (worst case:
 * density * (density/2)^2 lines of 2 * (integration_steps/2) + 1 points).  Blocking. */
typedef struct fluid_streamline_params {
    int density;                        /* STREAMLINE_DENSITY            GUI/config.py:20 (30) */
    double proximity;                   /* STREAMLINE_PROXIMITY          :21 (2)               */
    int integration_steps;              /* INTEGRATION_STEPS             :22 (100)             */
    double step_size;                   /* INTEGRATION_STEP_SIZE         :23 (0.2)             */
    double velocity_change_threshold;   /* VELOCITY_CHANGE_THRESHOLD     :24 (0.1)             */
} fluid_streamline_params;
int fluid_streamlines(fluid_handle*, const fluid_streamline_params* params, int max_lines, long long max_points,
                      int* n_lines, long long* n_points, int* line_len, double* norm_speed, double* points);

/* GUI/utils.py:10-38 (generate_obstacle_mesh): marching cubes of the padded obstacle volume at `level` (the viewer:
 * 0.5), vertices in padded (x, y, z) index coordinates as main_window.py:204 hands the volume over.  One vertex per
 * grid edge whose ends lie on different sides of the level (linear interpolation: the midpoint for a 0/1 volume),
 * faces as vertex-index triples with normals towards lower values (skimage's gradient_direction='descent').
 * *n_verts / *n_faces always receive the counts.  verts == faces == NULL: sizing query, returns 0.  Otherwise
 * verts (3 floats per vertex) and faces (3 ints per face) are filled when the counts fit max_verts / max_faces,
 * else nothing is copied and the call fails with FLUID_E_ARG.  No obstacle: 0 vertices, 0 faces (the viewer's
 * empty mesh).  The face list and the vertex order are this library's own: skimage's Lewiner variant is not
 * available here to pin them against. */
int fluid_obstacle_mesh(fluid_handle*, double level, long long max_verts, long long max_faces, long long* n_verts,
                        long long* n_faces, float* verts, int* faces);

int fluid_sync(fluid_handle*);
const char* fluid_last_error(fluid_handle*);

/* pinned host memory helpers for frame buffers (cudaHostAlloc / cudaFreeHost) */
void* fluid_host_alloc(size_t bytes);
void  fluid_host_free(void*);

/* device-side timing of the handle's stream (CUDA events): start, stop -> milliseconds */
int fluid_timer_start(fluid_handle*);
int fluid_timer_stop(fluid_handle*, float* ms);

/* per-solve device timing for bench.py's roofline: while enabled every linearSolver call
 * (simulation.cpp:251) is bracketed by CUDA events on the handle's stream.  read() blocks and
 * returns { total ms inside linearSolver, field-sweeps executed (fields x acc), solver calls }
 * since the last read, then resets. */
int fluid_profile_enable(fluid_handle*, int on);
int fluid_profile_read(fluid_handle*, double out[3]);

/* number of kernels this handle has launched so far (bench.py's gpu_launches) */
long long fluid_launch_count(fluid_handle*);

/* library/build info: "libfluid_cuda sm_100a ..." */
const char* fluid_version(void);

/* select the CUDA device subsequent fluid_create calls of this thread use (cudaSetDevice) */
int fluid_set_device(int device);

/* ---- multi-GPU z-slabs: one process per GPU on one NVLink/NVSwitch box ------------------------
 * Halo planes (two per interior side) travel by ncclSend/ncclRecv on the solver stream after every
 * sweep and around divergence / gradient / advect; libnccl is dlopen'ed from `nccl_lib_path` (NULL or
 * "" = the process's libnccl.so.2).  Rank 0 calls fluid_comm_unique_id and the host program hands the
 * 128 bytes to every rank (e.g. torch.distributed broadcast); fluid_comm_init is collective. */
int fluid_comm_unique_id(const char* nccl_lib_path, char out[128]);
int fluid_comm_init(fluid_handle*, const char* nccl_lib_path, const char unique_id[128]);
/* advect() gathers at arbitrary z (simulation.cpp:384-397): the gather sources v_x_prev, v_y_prev,
 * v_z_prev and buffer of every other rank are mapped through CUDA IPC (64-byte handles, exchanged by
 * the host program) and read over NVLink by the advection kernel.  Map v_x, v_y, v_z of every other rank as
 * well: step() then trades v and v_prev instead of copying (simulation.cpp:108-110), which makes those
 * allocations gather sources every other step; without them the three copies stay.
 * The linearSolver sweeps (simulation.cpp:258-272) hand their halo planes to the z-neighbours the same
 * way once the neighbours' allocations are mapped: ids 0..FLUID_NFIELDS-1 are the fields,
 * FLUID_IPC_SCRATCH0..+3 the solver's ping-pong partners and FLUID_IPC_SYNC the flag block (arrival flags
 * of the two z-neighbours, barrier epochs of every rank: import it from ALL ranks).  With all
 * of them imported from both neighbours a sweep ends in one small kernel that stores the two boundary
 * planes straight into the neighbour's halo and raises its flag (the neighbour's streams wait on the
 * flag with a stream memory operation); otherwise the exchange after a sweep is an ncclSend/ncclRecv. */
enum { FLUID_IPC_SCRATCH0 = FLUID_NFIELDS, FLUID_IPC_SYNC = FLUID_NFIELDS + 4, FLUID_IPC_COUNT = FLUID_NFIELDS + 5 };
int fluid_ipc_export(fluid_handle*, int alloc_id, char out[64]);
int fluid_ipc_import(fluid_handle*, int peer_rank, int alloc_id, const char in[64]);
/* device-side barrier across all ranks on the handle's stream (collective): epoch flags stored into every
 * rank's flag block over NVLink where those are mapped, a one-float ncclAllReduce otherwise */
int fluid_barrier(fluid_handle*);

#ifdef __cplusplus
}
#endif
#endif /* FLUID_CUDA_H */
