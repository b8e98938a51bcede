// simulation.cpp -- host side of the drop-in: the reference's Simulation API forwarded to libfluid_cuda.
#include "simulation.h"
#include "slab_hub.h"
#include "../../include/fluid_cuda.h"

#include <cstdlib>
#include <fcntl.h>

#include <iostream>
#include <stdexcept>

void Simulation::check(int rc, const char* what) const
{
    if (rc != 0)
        throw std::runtime_error(std::string(what) + ": " + fluid_last_error(h_));
}

Simulation::Simulation(int w, int h, int d, int iter_, int speed_, float dt_, float diff_, float visc_, int acc_,
                       unsigned flags_)
    : width(w), height(h), depth(d), speed(speed_), dt(dt_), diff(diff_), visc(visc_), acc(acc_), iter(iter_)
{
    flags = flags_;
    h_ = fluid_create(w, h, d, speed, dt, diff, visc, acc, flags);          // simulation.cpp:17-44
    if (!h_) throw std::runtime_error(std::string("fluid_create: ") + fluid_last_error(nullptr));
    d_local_ = d;
}

// z-slab member: the C++ counterpart of fluid_simulation_b200/multi_gpu.py:attach
Simulation::Simulation(slab::Hub& hub, int w, int h, int d, int iter_, int speed_, float dt_, float diff_, float visc_,
                       int acc_, unsigned flags_)
    : width(w), height(h), depth(d), speed(speed_), dt(dt_), diff(diff_), visc(visc_), acc(acc_), iter(iter_)
{
    flags = flags_; hub_ = &hub; rank_ = hub.rank; nranks_ = hub.nranks;
    h_ = fluid_create_slab(w, h, d, speed, dt, diff, visc, acc, flags, rank_, nranks_);
    if (!h_) throw std::runtime_error(std::string("fluid_create_slab: ") + fluid_last_error(nullptr));
    check(fluid_get_dims(h_, nullptr, nullptr, nullptr, &z_lo_, &d_local_), "get_dims");
    if (nranks_ == 1) return;
    const char* lib = std::getenv("FLUID_NCCL_LIB");                        // default: the process's libnccl.so.2
    // the 128-byte NCCL id of rank 0, then the collective communicator set-up
    std::vector<char> ids((size_t)128 * nranks_), uid(128, 0);
    if (rank_ == 0) check(fluid_comm_unique_id(lib ? lib : "", uid.data()), "comm_unique_id");
    hub.allgather(uid.data(), 128, ids.data());
    check(fluid_comm_init(h_, lib ? lib : "", ids.data()), "comm_init");
    // CUDA IPC: every solver buffer and the flag block of the two z-neighbours (halos are handed over by peer
    // stores), the advection gather sources and the flag block of every other rank
    const size_t per = (size_t)FLUID_IPC_COUNT * 64;
    std::vector<char> mine(per), all(per * nranks_);
    for (int a = 0; a < FLUID_IPC_COUNT; ++a) check(fluid_ipc_export(h_, a, mine.data() + 64 * a), "ipc_export");
    hub.allgather(mine.data(), per, all.data());
    // (v_x, v_y, v_z too: step() trades them with the *_prev buffers instead of copying, so they are gather sources every other step)
    static const int gather_ids[8] = { FLUID_VX_PREV, FLUID_VY_PREV, FLUID_VZ_PREV, FLUID_BUFFER, FLUID_VX, FLUID_VY, FLUID_VZ, FLUID_IPC_SYNC };
    for (int r = 0; r < nranks_; ++r) {
        if (r == rank_) continue;
        const char* hr = all.data() + per * r;
        if (r == rank_ - 1 || r == rank_ + 1) { for (int a = 0; a < FLUID_IPC_COUNT; ++a) check(fluid_ipc_import(h_, r, a, hr + 64 * a), "ipc_import"); }
        else for (int a : gather_ids) check(fluid_ipc_import(h_, r, a, hr + 64 * a), "ipc_import");
    }
    hub.barrier();
}

// speed, dt, diff, visc and acc are public and mutable like the reference's (simulation.h:44-54)
void Simulation::push_params() { check(fluid_set_params(h_, speed, dt, diff, visc, acc), "set_params"); }

// run()'s statistics over the whole tunnel: per-slab partials folded over the ranks (sum adds, min / max fold)
void Simulation::reduced_statistics(double out[9])
{
    statistics(out);
    if (nranks_ == 1) return;
    std::vector<double> all((size_t)9 * nranks_);
    hub_->allgather(out, 9 * sizeof(double), all.data());
    for (int k = 0; k < 9; ++k) {
        double v = all[k];
        for (int r = 1; r < nranks_; ++r) {
            const double x = all[(size_t)9 * r + k];
            v = k == 0 ? v + x : (k % 2 == 1 ? (x < v ? x : v) : (x > v ? x : v));      // [0] sum; odd: minima; even: maxima
        }
        out[k] = v;
    }
}

Simulation::~Simulation()
{
    for (auto& f : files_) if (f) std::fclose(f);
    for (auto& p : frame_) if (p) fluid_host_free(p);
    fluid_destroy(h_);
}

void Simulation::stage(Kind k)
{
    if (pending_ != NONE && pending_ != k) flush();     // keep the caller's order between kinds
    pending_ = k;
}

void Simulation::flush()
{
    const int n = (int)(xyz_.size() / 3);
    if (n) {
        if (pending_ == OBS) check(fluid_add_obstacles(h_, n, xyz_.data()), "addObstacle");
        if (pending_ == DENS) check(fluid_add_density(h_, n, xyz_.data(), val_.data()), "addDensity");
        if (pending_ == VEL) check(fluid_set_velocity(h_, n, xyz_.data(), val_.data()), "setVelocity");
        check(fluid_sync(h_), "flush");                 // the staging vectors are reused
    }
    xyz_.clear(); val_.clear(); pending_ = NONE;
}

void Simulation::addObstacle(int x, int y, int z)
{
    stage(OBS);
    xyz_.insert(xyz_.end(), { x, y, z });
}

void Simulation::addDensity(int x, int y, int z, float amount)
{
    stage(DENS);
    xyz_.insert(xyz_.end(), { x, y, z }); val_.push_back(amount);
}

void Simulation::setVelocity(int x, int y, int z, float ax, float ay, float az)
{
    stage(VEL);
    xyz_.insert(xyz_.end(), { x, y, z }); val_.insert(val_.end(), { ax, ay, az });
}

void Simulation::uploadObstacles(const float* obs_padded)
{
    flush();
    check(fluid_upload_obstacles(h_, obs_padded), "uploadObstacles");
}

void Simulation::downloadField(int field_id, float* padded_host)
{
    flush();
    check(fluid_download_field(h_, field_id, padded_host), "downloadField");
}

long Simulation::restartFromFrames(const std::string& dir, long frame)
{
    // a dumped frame holds dens, obs and v; the other six fields are scratch that step() overwrites before
    // reading them (simulation.cpp:96-150), so nothing else is needed
    static const char* names[5] = { "data.bin", "obs.bin", "v_x.bin", "v_y.bin", "v_z.bin" };
    static const int ids[5] = { FLUID_DENS, FLUID_OBS, FLUID_VX, FLUID_VY, FLUID_VZ };
    const size_t npad = (size_t)(width + 2) * (height + 2) * (depth + 2);
    std::vector<float> buf(npad);
    long frames = -1;
    std::FILE* fs[5] = {};
    auto close_all = [&]() { for (auto& f : fs) if (f) { std::fclose(f); f = nullptr; } };
    for (int i = 0; i < 5; ++i) {
        fs[i] = std::fopen((dir + "/" + names[i]).c_str(), "rb");
        if (!fs[i]) { close_all(); throw std::runtime_error("restart: cannot open " + dir + "/" + names[i]); }
        std::fseek(fs[i], 0, SEEK_END);
        const long n = (long)((size_t)std::ftell(fs[i]) / (npad * sizeof(float)));
        frames = frames < 0 || n < frames ? n : frames;
    }
    const long k = frame >= 0 ? frame : frames + frame;
    if (frames < 1 || k < 0 || k >= frames) { close_all(); throw std::runtime_error("restart: frame out of range in " + dir); }
    flush();
    for (int i = 0; i < 5; ++i) {
        std::fseek(fs[i], (long)((size_t)k * npad * sizeof(float)), SEEK_SET);
        if (std::fread(buf.data(), sizeof(float), npad, fs[i]) != npad) { close_all(); throw std::runtime_error("restart: short read"); }
        if (ids[i] == FLUID_OBS) check(fluid_upload_obstacles(h_, buf.data()), "restart obs");
        else check(fluid_upload_field(h_, ids[i], buf.data()), "restart field");
    }
    close_all();
    return k;
}

void Simulation::statistics(double out[9])
{
    flush();
    check(fluid_stats(h_, out), "statistics");
}

void Simulation::write_pending_frame()
{
    if (!frame_pending_) return;
    check(fluid_frame_wait(h_), "frame wait");
    const size_t plane = (size_t)(width + 2) * (height + 2);
    if (nranks_ == 1) {
        for (int i = 0; i < 5; ++i)
            if (files_[i]) std::fwrite(frame_[i], sizeof(float), plane * (depth + 2), files_[i]);  // simulation.cpp:143-147
    } else {
        // z-slabs share the files: this rank's planes go to their place inside the frame
        const off_t off = (off_t)((frames_written_ * (depth + 2) + own_first()) * plane * sizeof(float));
        for (int i = 0; i < 5; ++i)
            if (files_[i] && ::pwrite(fileno(files_[i]), frame_[i], own_count() * plane * sizeof(float), off) < 0)
                throw std::runtime_error("frame write failed");
    }
    ++frames_written_;
    frame_pending_ = false;
}

// One step as Simulation::step() does it: the solver phases, then the frame write.  The D2H copy runs
// on the library's side stream; the bytes reach the files while the NEXT step computes.
void Simulation::step()
{
    flush();
    push_params();
    check(fluid_call(h_, FLUID_OP_STEP, 0, 0, 0, 0.0f, 0.0f), "step");
    bool any = false;
    for (auto& f : files_) any = any || f;
    if (!any) return;                       // like the reference with unopened streams: no frame I/O
    write_pending_frame();
    // pinned buffers for the planes this rank owns; the C ABI takes the base of a whole-frame buffer
    const size_t plane = (size_t)(width + 2) * (height + 2);
    const size_t bytes = plane * own_count() * sizeof(float);
    float* base[5];
    for (int i = 0; i < 5; ++i) {
        if (!frame_[i] && !(frame_[i] = (float*)fluid_host_alloc(bytes))) throw std::runtime_error("pinned frame buffer allocation failed");
        base[i] = frame_[i] - plane * own_first();
    }
    check(fluid_dump_frame_async(h_, base), "frame dump");
    frame_pending_ = true;
}

void Simulation::run()
{
    const bool lead = rank_ == 0;
    const bool talk = !quiet && lead;
    if (talk)
        std::cout << "starting 3-D simulation: " << width << "x" << height << "x" << depth
                  << "  steps = " << iter << (nranks_ > 1 ? "  (z-slabs on " + std::to_string(nranks_) + " GPUs)" : std::string()) << "\n";
    // simulation.cpp:56-60; a missing directory silently disables the output, like the reference's ofstreams.
    // z-slabs: rank 0 creates and truncates the files, the others open them once they exist.
    static const char* names[5] = { "data.bin", "obs.bin", "v_x.bin", "v_y.bin", "v_z.bin" };
    frames_written_ = 0;
    if (lead) for (int i = 0; i < 5; ++i) files_[i] = std::fopen((out_dir_ + "/" + names[i]).c_str(), "wb");
    if (nranks_ > 1) {
        hub_->barrier();
        if (!lead) for (int i = 0; i < 5; ++i) files_[i] = std::fopen((out_dir_ + "/" + names[i]).c_str(), "r+b");
    }
    flush();
    push_params();
    check(fluid_sync(h_), "sync");
    check(fluid_timer_start(h_), "timer");
    for (int i = 0; i < iter; ++i) {
        check(fluid_call(h_, FLUID_OP_INFLOW_DENS, 0, 0, 0, 0.0f, 0.0f), "inlet density");   // :65-67
        check(fluid_call(h_, FLUID_OP_COPY, 0, FLUID_BUFFER, FLUID_DENS, 0.0f, 0.0f), "buffer = dens"); // :70
        if ((i + 1) % dump_every == 0) step();                                                // :71
        else check(fluid_call(h_, FLUID_OP_STEP, 0, 0, 0, 0.0f, 0.0f), "step");
        if ((i + 1) % 100 == 0 && i > 0 && !quiet) {                                          // :73-77
            double st[9]; reduced_statistics(st);                                            // collective on z-slabs
            if (lead) std::cout << "step " << i + 1 << "\n  density sum = " << st[0] << "\n";
        }
    }
    float ms = 0.0f;
    check(fluid_timer_stop(h_, &ms), "timer");
    ms_per_step_ = iter > 0 ? ms / iter : 0.0;
    write_pending_frame();
    const bool wrote = files_[0] != nullptr;
    for (auto& f : files_) if (f) { std::fclose(f); f = nullptr; }
    if (nranks_ > 1) hub_->barrier();                   // every rank's planes are in the files
    if (wrote && lead) {
        // what the reference's viewers keep in hand-synchronised constants (GUI/config.py:8-11, gui.py:32-34)
        if (std::FILE* m = std::fopen((out_dir_ + "/meta.json").c_str(), "w")) {
            std::fprintf(m, "{\n \"width\": %d, \"height\": %d, \"depth\": %d, \"padded\": [%d, %d, %d],\n \"frames\": %d, \"dump_every\": %d,"
                            " \"dtype\": \"float32\", \"order\": \"z,y,x (C order, padded)\",\n \"files\": [\"data.bin\", \"obs.bin\", \"v_x.bin\", \"v_y.bin\", \"v_z.bin\"],\n"
                            " \"speed\": %d, \"dt\": %.9g, \"diff\": %.9g, \"visc\": %.9g, \"acc\": %d\n}\n",
                         width, height, depth, depth + 2, height + 2, width + 2, iter / dump_every, dump_every, speed, dt, diff, visc, acc);
            std::fclose(m);
        }
    }
    double st[9]; reduced_statistics(st);
    if (talk) {                                                                               // :81-90
        std::cout << "\n--- statistics -------------------------------------------------\n";
        std::cout << "density  min = " << st[1] << "\n" << "density  max = " << st[2] << "\n";
        std::cout << "velocity x min = " << st[3] << "\n" << "velocity x max = " << st[4] << "\n";
        std::cout << "velocity y min = " << st[5] << "\n" << "velocity y max = " << st[6] << "\n";
        std::cout << "velocity z min = " << st[7] << "\n" << "velocity z max = " << st[8] << "\n";
        std::cout << "simulation finished\n";
    }
}
