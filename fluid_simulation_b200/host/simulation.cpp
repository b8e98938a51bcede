// simulation.cpp -- host side of the drop-in: the reference's Simulation API forwarded to libfluid_cuda.
#include "simulation.h"
#include "../../include/fluid_cuda.h"

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
    const size_t n = (size_t)(width + 2) * (height + 2) * (depth + 2);
    for (int i = 0; i < 5; ++i)
        if (files_[i]) std::fwrite(frame_[i], sizeof(float), n, files_[i]);  // simulation.cpp:143-147
    frame_pending_ = false;
}

// One step as Simulation::step() does it: the solver phases, then the frame write.  The D2H copy runs
// on the library's side stream; the bytes reach the files while the NEXT step computes.
void Simulation::step()
{
    flush();
    check(fluid_call(h_, FLUID_OP_STEP, 0, 0, 0, 0.0f, 0.0f), "step");
    bool any = false;
    for (auto& f : files_) any = any || f;
    if (!any) return;                       // like the reference with unopened streams: no frame I/O
    write_pending_frame();
    const size_t bytes = (size_t)(width + 2) * (height + 2) * (depth + 2) * sizeof(float);
    for (auto& p : frame_)
        if (!p && !(p = (float*)fluid_host_alloc(bytes))) throw std::runtime_error("pinned frame buffer allocation failed");
    check(fluid_dump_frame_async(h_, frame_), "frame dump");
    frame_pending_ = true;
}

void Simulation::run()
{
    if (!quiet)
        std::cout << "starting 3-D simulation: " << width << "x" << height << "x" << depth
                  << "  steps = " << iter << "\n";
    // simulation.cpp:56-60; a missing directory silently disables the output, like the reference's ofstreams
    static const char* names[5] = { "data.bin", "obs.bin", "v_x.bin", "v_y.bin", "v_z.bin" };
    for (int i = 0; i < 5; ++i) files_[i] = std::fopen((out_dir_ + "/" + names[i]).c_str(), "wb");
    flush();
    check(fluid_sync(h_), "sync");
    check(fluid_timer_start(h_), "timer");
    for (int i = 0; i < iter; ++i) {
        check(fluid_call(h_, FLUID_OP_INFLOW_DENS, 0, 0, 0, 0.0f, 0.0f), "inlet density");   // :65-67
        check(fluid_call(h_, FLUID_OP_COPY, 0, FLUID_BUFFER, FLUID_DENS, 0.0f, 0.0f), "buffer = dens"); // :70
        if ((i + 1) % dump_every == 0) step();                                                // :71
        else check(fluid_call(h_, FLUID_OP_STEP, 0, 0, 0, 0.0f, 0.0f), "step");
        if ((i + 1) % 100 == 0 && i > 0 && !quiet) {                                          // :73-77
            double st[9]; statistics(st);
            std::cout << "step " << i + 1 << "\n  density sum = " << st[0] << "\n";
        }
    }
    float ms = 0.0f;
    check(fluid_timer_stop(h_, &ms), "timer");
    ms_per_step_ = iter > 0 ? ms / iter : 0.0;
    write_pending_frame();
    const bool wrote = files_[0] != nullptr;
    for (auto& f : files_) if (f) { std::fclose(f); f = nullptr; }
    if (wrote) {
        // what the reference's viewers keep in hand-synchronised constants (GUI/config.py:8-11, gui.py:32-34)
        if (std::FILE* m = std::fopen((out_dir_ + "/meta.json").c_str(), "w")) {
            std::fprintf(m, "{\n \"width\": %d, \"height\": %d, \"depth\": %d, \"padded\": [%d, %d, %d],\n \"frames\": %d, \"dump_every\": %d,"
                            " \"dtype\": \"float32\", \"order\": \"z,y,x (C order, padded)\",\n \"files\": [\"data.bin\", \"obs.bin\", \"v_x.bin\", \"v_y.bin\", \"v_z.bin\"],\n"
                            " \"speed\": %d, \"dt\": %.9g, \"diff\": %.9g, \"visc\": %.9g, \"acc\": %d\n}\n",
                         width, height, depth, depth + 2, height + 2, width + 2, iter / dump_every, dump_every, speed, dt, diff, visc, acc);
            std::fclose(m);
        }
    }
    double st[9]; statistics(st);
    if (!quiet) {                                                                             // :81-90
        std::cout << "\n--- statistics -------------------------------------------------\n";
        std::cout << "density  min = " << st[1] << "\n" << "density  max = " << st[2] << "\n";
        std::cout << "velocity x min = " << st[3] << "\n" << "velocity x max = " << st[4] << "\n";
        std::cout << "velocity y min = " << st[5] << "\n" << "velocity y max = " << st[6] << "\n";
        std::cout << "velocity z min = " << st[7] << "\n" << "velocity z max = " << st[8] << "\n";
        std::cout << "simulation finished\n";
    }
}
