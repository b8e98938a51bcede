// object_loader.cpp -- host side of the STL loader: parse the file, hand the triangles to the device.
// Parsing follows the acceptance rules of the reference reader (object_loader.cpp:98-174): a file is
// binary unless its first line, stripped, begins with "solid"; binary facets are 50-byte records after an
// 80-byte header and a uint32 count; ASCII facets are kept when exactly three "vertex" lines were seen
// between "outer loop" and "endfacet".
#include "object_loader.h"
#include "simulation.h"
#include "../../include/fluid_cuda.h"

#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>

namespace {

std::string strip(const std::string& s)
{
    const char* blanks = " \t\n\r";
    const size_t a = s.find_first_not_of(blanks);
    if (a == std::string::npos) return std::string();
    return s.substr(a, s.find_last_not_of(blanks) - a + 1);
}

std::vector<float> parse_binary(const std::string& raw)
{
    std::vector<float> out;
    if (raw.size() < 84) return out;
    uint32_t n = 0;
    std::memcpy(&n, raw.data() + 80, 4);
    const size_t have = (raw.size() - 84) / 50;
    if (n > have) n = (uint32_t)have;                       // truncated file: keep the complete records
    out.resize((size_t)n * 9);
    for (uint32_t i = 0; i < n; ++i)                        // 12 bytes normal, 36 bytes vertices, 2 bytes attribute
        std::memcpy(&out[(size_t)i * 9], raw.data() + 84 + (size_t)i * 50 + 12, 36);
    return out;
}

std::vector<float> parse_ascii(const std::string& raw)
{
    std::vector<float> out;
    std::istringstream in(raw);
    std::string line;
    std::getline(in, line);                                 // "solid ..."
    float tri[9] = {};
    int seen = 0;
    while (std::getline(in, line)) {
        line = strip(line);
        if (line == "outer loop") { seen = 0; continue; }
        if (line == "endfacet") { if (seen == 3) out.insert(out.end(), tri, tri + 9); continue; }
        if (line.compare(0, 6, "vertex") != 0) continue;
        std::istringstream v(line.substr(6));
        float x, y, z;
        if (!(v >> x >> y >> z)) continue;
        if (seen < 3) { tri[3 * seen] = x; tri[3 * seen + 1] = y; tri[3 * seen + 2] = z; }
        seen = (seen + 1) % 4;                              // a fourth vertex wraps the count: facet dropped
    }
    return out;
}

}  // namespace

std::vector<float> readSTLTriangles(const std::string& filename)
{
    std::ifstream f(filename, std::ios::binary);
    if (!f) {
        std::cerr << "Error: Cannot open STL file: " << filename << "\n";
        return {};
    }
    std::ostringstream buf;
    buf << f.rdbuf();
    const std::string raw = buf.str();
    const std::string first = strip(raw.substr(0, raw.find('\n')));
    return first.compare(0, 5, "solid") == 0 ? parse_ascii(raw) : parse_binary(raw);
}

long long loadSTLIntoObstacles(const char* stlFile, Simulation& sim, int sampling, float scale, float rot_x,
                               float rot_y, float rot_z, float translate_x, float translate_y, float translate_z)
{
    const std::vector<float> tris = readSTLTriangles(stlFile);
    if (!sim.quiet) std::cout << "Loaded " << tris.size() / 9 << " triangles.\n";
    if (tris.empty()) {
        std::cerr << "Failed to load STL: " << stlFile << "\n";
        return 0;
    }
    long long added = 0;
    fluid_handle* h = sim.handle();                         // flushes staged addObstacle / addDensity calls first
    const int rc = fluid_voxelize(h, tris.data(), (int)(tris.size() / 9), scale, rot_x, rot_y, rot_z, translate_x,
                                  translate_y, translate_z, sampling, &added);
    if (rc != 0) throw std::runtime_error(std::string("loadSTLIntoObstacles: ") + fluid_last_error(h));
    if (!sim.quiet) std::cout << "Added " << added << " obstacle points.\n";
    return added;
}

void loadSTLIntoObstacles(const char* stlFile, Simulation& sim, float scale, float rot_x, float rot_y, float rot_z,
                          float translate_x, float translate_y, float translate_z)
{
    loadSTLIntoObstacles(stlFile, sim, FLUID_VOX_HASHED, scale, rot_x, rot_y, rot_z, translate_x, translate_y, translate_z);
}
