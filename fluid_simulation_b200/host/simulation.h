// simulation.h -- the reference's `Simulation` class (/root/reference/simulation.h:11-91) over
// libfluid_cuda: same public data members, constructor and methods; the eleven std::vector<float>
// fields and the five private solver methods are replaced by a fluid_handle* (include/fluid_cuda.h).
// All compute runs on the B200; there is no CPU path (the constructor throws if no device is usable).
#ifndef FLUID_B200_SIMULATION_H
#define FLUID_B200_SIMULATION_H

#include <cstdio>
#include <string>
#include <vector>

struct fluid_handle;
namespace slab { class Hub; }

class Simulation {
private:
    fluid_handle* h_ = nullptr;

    // addObstacle / addDensity / setVelocity are per-cell calls (the STL loader makes millions,
    // object_loader.cpp:439-442): they are staged here and flushed in order before the next device op
    enum Kind { NONE, OBS, DENS, VEL };
    Kind pending_ = NONE;
    std::vector<int> xyz_;
    std::vector<float> val_;
    void stage(Kind k);
    void flush();

    // frame output (simulation.cpp:56-60, 140-148): pinned double buffer, written while the next step runs
    std::string out_dir_ = "data";
    std::FILE* files_[5] = {};
    float* frame_[5] = {};
    bool frame_pending_ = false;
    long frames_written_ = 0;
    void write_pending_frame();
    void check(int rc, const char* what) const;

public:
    /* grid dimensions (without padding), inlet speed, parameters: simulation.h:44-54 */
    int width, height, depth;
    int speed;
    float dt, diff, visc;
    int acc, iter;

    /* extras of the B200 build.  flags = 0 selects the red-black sweep order (FLUID_SOLVER_RB), which is NOT bit-identical
     * to the reference's single-thread lexicographic Gauss-Seidel; FLUID_SOLVER_LEX is the bit-exact mode. */
    unsigned flags = 0;          // FLUID_SOLVER_RB (default) | FLUID_SOLVER_JACOBI | FLUID_SOLVER_LEX | FLUID_ELIDE_DEAD
    int dump_every = 1;          // the reference writes a frame every step (simulation.cpp:140-148)
    bool quiet = false;

    Simulation(int w, int h, int d, int iter, int speed = 30, float dt = 0.05f, float diff = 2.0e-5f,
               float visc = 1.5e-5f, int acc = 15, unsigned flags = 0);
    /* one z-slab of a multi-GPU run (one process per GPU, host/slab_hub.h): collective over the hub's ranks.
     * Creates the slab handle on the process's current device, attaches the NCCL communicator and maps the peers'
     * buffers; run() then writes this rank's planes into the shared frame files and reduces the statistics. */
    Simulation(slab::Hub& hub, int w, int h, int d, int iter, int speed = 30, float dt = 0.05f, float diff = 2.0e-5f,
               float visc = 1.5e-5f, int acc = 15, unsigned flags = 0);
    ~Simulation();
    Simulation(const Simulation&) = delete;
    Simulation& operator=(const Simulation&) = delete;

    void run();                                   // simulation.cpp:49-91
    void step();                                  // simulation.cpp:96-150 (incl. the frame write)
    void addObstacle(int x, int y, int z);        // simulation.cpp:155
    void addDensity(int x, int y, int z, float amount);                       // :163
    void setVelocity(int x, int y, int z, float ax, float ay, float az);      // :172

    /* access that the reference gave through its (private) vectors */
    void setOutputDirectory(const std::string& dir) { out_dir_ = dir; }
    void uploadObstacles(const float* obs_padded);                            // whole mask at once
    void downloadField(int field_id, float* padded_host);
    /* restart: load frame `frame` (negative = from the end) of dir/{data,obs,v_x,v_y,v_z}.bin as
     * the current state; the following steps reproduce the original run bit for bit.  Returns the frame index. */
    long restartFromFrames(const std::string& dir, long frame = -1);
    void statistics(double out[9]);
    double msPerStepLastRun() const { return ms_per_step_; }
    fluid_handle* handle() { flush(); return h_; }

private:
    double ms_per_step_ = 0.0;
    slab::Hub* hub_ = nullptr;   // z-slab member: rendezvous with the other ranks
    int rank_ = 0, nranks_ = 1, z_lo_ = 1, d_local_ = 0;
    int own_first() const { return rank_ == 0 ? 0 : z_lo_; }                                   // padded planes this rank owns
    int own_count() const { return (rank_ == nranks_ - 1 ? depth + 1 : z_lo_ + d_local_ - 1) - own_first() + 1; }
    void push_params();
    void reduced_statistics(double out[9]);
};

/* loadSTLIntoObstacles (the reference's object_loader.h) lives in object_loader.h: the mesh is voxelised on
 * the device.  The reference's own object_loader.cpp also still compiles against this class (it only uses
 * width / height / depth and addObstacle). */

#endif
