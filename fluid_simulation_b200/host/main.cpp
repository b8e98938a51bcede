// fluid_tunnel -- parametrised replacement of the reference's hard-coded main() (simulation.cpp:429-450):
// builds a Simulation, installs an obstacle mask, runs `iter` steps, writes data/*.bin frames in the
// reference's format and prints run()'s statistics plus the device time per step.
//   fluid_tunnel [--w 128 --h 64 --d 64 --iter 100 --acc 15 --speed 30] [--obstacle none|sphere|box|FILE.bin|FILE.stl]
//                [--restart DIR[:FRAME]] [--stl-scale 0.8 --stl-rot rx,ry,rz --stl-translate tx,ty,tz --stl-sampling hashed|fixed|cells]
//                [--solver rb|jacobi|lex] [--elide] [--out data] [--dump-every 1] [--quiet] [--gpus N]
//                [--view-ppm FILE [--view-field dens|v_x|v_y|v_z] [--view-z N]] [--view-streamlines FILE]
// --view-*: after the run, what the reference's viewers would draw from the last frame (gui.py:265-300 z-slice as a
// binary PPM; GUI/utils.py:118-214 streamlines as text: "line <n> <norm_speed>" followed by n "x y z" rows, origin
// shifted by -1 as GUI/main_window.py:236 does), computed on the device.
// --gpus N (N > 1): z-slabs on N GPUs of this node -- one forked process per GPU (slab_hub.h), halos handed over between
// the GPUs inside libfluid_cuda, frames written by every rank into the same files; the output equals the 1-GPU run's.
// FILE.bin = raw float32 padded mask (w+2)(h+2)(d+2) in the reference layout, e.g. dumped once from the
// reference's STL loader (its output is not reproducible run to run).
#include "simulation.h"
#include "slab_hub.h"
#include "object_loader.h"
#include "../../include/fluid_cuda.h"

#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <stdexcept>

static std::vector<float> analytic_mask(const std::string& kind, int w, int h, int d)
{
    std::vector<float> m((size_t)(w + 2) * (h + 2) * (d + 2), 0.0f);
    auto at = [&](int x, int y, int z) -> float& { return m[(size_t)x + (size_t)y * (w + 2) + (size_t)z * (w + 2) * (h + 2)]; };
    for (int z = 1; z <= d; ++z)
        for (int y = 1; y <= h; ++y)
            for (int x = 1; x <= w; ++x) {
                bool in = false;
                if (kind == "sphere") {          // dist to (0.35W, 0.5H, 0.5D) <= 0.15H, float32
                    const float cx = (float)(0.35 * w), cy = (float)(0.5 * h), cz = (float)(0.5 * d), r = (float)(0.15 * h);
                    const float dx = (float)x - cx, dy = (float)y - cy, dz = (float)z - cz;
                    in = dx * dx + dy * dy + dz * dz <= r * r;
                } else if (kind == "box") {
                    in = x >= 0.3 * w && x <= 0.45 * w && y >= 0.35 * h && y <= 0.65 * h && z >= 0.35 * d && z <= 0.65 * d;
                }
                if (in) at(x, y, z) = 1.0f;
            }
    return m;
}

int main(int argc, char** argv)
{
    int w = 128, h = 64, d = 64, iter = 100, acc = 15, speed = 30, dump_every = 1;   // simulation.cpp:431-436
    std::string obstacle = "none", solver = "rb", out = "data", restart, view_ppm, view_lines, view_field = "dens";
    int view_z = -1, gpus = 1;
    bool elide = false, quiet = false;
    float stl_scale = 0.8f, stl_rot[3] = { 0, 0, 0 }, stl_tr[3] = { 0, 0, 0 };      // loadSTLIntoObstacles defaults
    int stl_sampling = FLUID_VOX_HASHED;
    auto triple = [](const char* s, float out[3]) { if (std::sscanf(s, "%f,%f,%f", out, out + 1, out + 2) != 3) { std::cerr << "expected a,b,c: " << s << "\n"; std::exit(2); } };
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto val = [&]() -> const char* { if (i + 1 >= argc) { std::cerr << a << " needs a value\n"; std::exit(2); } return argv[++i]; };
        if (a == "--w") w = std::atoi(val()); else if (a == "--h") h = std::atoi(val());
        else if (a == "--d") d = std::atoi(val()); else if (a == "--iter") iter = std::atoi(val());
        else if (a == "--acc") acc = std::atoi(val()); else if (a == "--speed") speed = std::atoi(val());
        else if (a == "--obstacle") obstacle = val(); else if (a == "--solver") solver = val();
        else if (a == "--out") out = val(); else if (a == "--dump-every") dump_every = std::atoi(val());
        else if (a == "--elide") elide = true; else if (a == "--quiet") quiet = true;
        else if (a == "--restart") restart = val(); else if (a == "--gpus") gpus = std::atoi(val());
        else if (a == "--view-ppm") view_ppm = val(); else if (a == "--view-streamlines") view_lines = val();
        else if (a == "--view-field") view_field = val(); else if (a == "--view-z") view_z = std::atoi(val());
        else if (a == "--stl-scale") stl_scale = (float)std::atof(val());
        else if (a == "--stl-rot") triple(val(), stl_rot); else if (a == "--stl-translate") triple(val(), stl_tr);
        else if (a == "--stl-sampling") { std::string m = val(); stl_sampling = m == "fixed" ? FLUID_VOX_FIXED : m == "cells" ? FLUID_VOX_CELLS : FLUID_VOX_HASHED; }
        else { std::cerr << "unknown argument " << a << "\n"; return 2; }
    }
    unsigned flags = solver == "jacobi" ? FLUID_SOLVER_JACOBI : solver == "lex" ? FLUID_SOLVER_LEX : FLUID_SOLVER_RB;
    if (elide) flags |= FLUID_ELIDE_DEAD;
    slab::Hub* hub = nullptr;
    if (gpus > 1) {
        if (!view_ppm.empty() || !view_lines.empty() || (obstacle.size() > 4 && obstacle.compare(obstacle.size() - 4, 4, ".stl") == 0)) {
            std::cerr << "fluid_tunnel: --view-* and STL voxelisation run on a single-GPU handle (omit --gpus)\n";
            return 2;
        }
        hub = slab::spawn(gpus);                        // the parent stays behind as the hub and exits with the ranks' status
        if (fluid_set_device(hub->rank) != 0) { std::cerr << "fluid_tunnel: rank " << hub->rank << ": " << fluid_last_error(nullptr) << "\n"; return 1; }
    }
    const bool lead = !hub || hub->rank == 0;
    try {
        std::unique_ptr<Simulation> simp(hub ? new Simulation(*hub, w, h, d, iter, speed, 0.05f, 2.0e-5f, 1.5e-5f, acc, flags)
                                             : new Simulation(w, h, d, iter, speed, 0.05f, 2.0e-5f, 1.5e-5f, acc, flags));
        Simulation& sim = *simp;
        sim.quiet = quiet; sim.dump_every = dump_every > 0 ? dump_every : 1;
        sim.setOutputDirectory(out);
        if (obstacle == "sphere" || obstacle == "box") {
            std::vector<float> m = analytic_mask(obstacle, w, h, d);
            sim.uploadObstacles(m.data());
        } else if (obstacle.size() > 4 && (obstacle.compare(obstacle.size() - 4, 4, ".stl") == 0 || obstacle.compare(obstacle.size() - 4, 4, ".STL") == 0)) {
            // the reference's main() route (simulation.cpp:441): voxelise a mesh, on the device
            loadSTLIntoObstacles(obstacle.c_str(), sim, stl_sampling, stl_scale, stl_rot[0], stl_rot[1], stl_rot[2], stl_tr[0], stl_tr[1], stl_tr[2]);
        } else if (obstacle != "none") {
            std::vector<float> m((size_t)(w + 2) * (h + 2) * (d + 2));
            std::ifstream f(obstacle, std::ios::binary);
            if (!f.read((char*)m.data(), (std::streamsize)(m.size() * sizeof(float))))
                throw std::runtime_error("cannot read obstacle mask " + obstacle);
            sim.uploadObstacles(m.data());
        }
        if (!restart.empty()) {                      // DIR or DIR:FRAME; the frame's obs replaces --obstacle
            const size_t c = restart.rfind(':');
            const long k = sim.restartFromFrames(c == std::string::npos ? restart : restart.substr(0, c),
                                                 c == std::string::npos ? -1 : std::atol(restart.c_str() + c + 1));
            if (!quiet && lead) std::cout << "restarted from frame " << k << "\n";
        }
        sim.run();
        if (!view_ppm.empty()) {                     // gui.py:265-300 with the viewer's fixed ranges (:272-290)
            const int fid = view_field == "v_x" ? FLUID_VX : view_field == "v_y" ? FLUID_VY : view_field == "v_z" ? FLUID_VZ : FLUID_DENS;
            const double lo = fid == FLUID_DENS ? 0.0 : fid == FLUID_VX ? -10.0 : -1.0, hi = fid == FLUID_DENS ? 0.01 : fid == FLUID_VX ? 10.0 : 1.0;
            std::vector<unsigned char> rgb((size_t)(w + 2) * (h + 2) * 3);
            if (fluid_render_slice(sim.handle(), fid, view_z < 0 ? (d + 2) / 2 : view_z, lo, hi, 0.2, rgb.data()) != 0)
                throw std::runtime_error(std::string("render_slice: ") + fluid_last_error(sim.handle()));
            std::ofstream f(view_ppm, std::ios::binary);
            f << "P6\n" << w + 2 << " " << h + 2 << "\n255\n";
            f.write((const char*)rgb.data(), (std::streamsize)rgb.size());
        }
        if (!view_lines.empty()) {                   // GUI/config.py:19-24 defaults
            const fluid_streamline_params sp = { 30, 2.0, 100, 0.2, 0.1 };
            const int max_lines = 30 * 15 * 15; const long long max_points = (long long)max_lines * 101;
            std::vector<int> len(max_lines); std::vector<double> norm(max_lines), pts((size_t)max_points * 3);
            int nl = 0; long long np = 0;
            if (fluid_streamlines(sim.handle(), &sp, max_lines, max_points, &nl, &np, len.data(), norm.data(), pts.data()) != 0)
                throw std::runtime_error(std::string("streamlines: ") + fluid_last_error(sim.handle()));
            std::FILE* f = std::fopen(view_lines.c_str(), "w");
            if (!f) throw std::runtime_error("cannot write " + view_lines);
            size_t k = 0;
            for (int i = 0; i < nl; ++i) {
                std::fprintf(f, "line %d %.17g\n", len[i], norm[i]);
                for (int j = 0; j < len[i]; ++j, ++k) std::fprintf(f, "%.17g %.17g %.17g\n", pts[3 * k] - 1, pts[3 * k + 1] - 1, pts[3 * k + 2] - 1);
            }
            std::fclose(f);
            if (!quiet) std::cout << nl << " streamlines, " << np << " points -> " << view_lines << "\n";
        }
        if (!quiet && lead) {
            const double cells = (double)w * h * d;
            std::cout << fluid_version() << "\n" << sim.msPerStepLastRun() << " ms per step (incl. frame dumps), "
                      << cells / (sim.msPerStepLastRun() * 1e-3) / 1e9 << " G cell-updates/s\n";
        }
    } catch (const std::exception& e) {
        std::cerr << "fluid_tunnel: " << e.what() << "\n";
        return 1;
    }
    return 0;
}
