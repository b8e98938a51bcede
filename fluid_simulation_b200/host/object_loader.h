// object_loader.h -- the reference's STL loader interface (/root/reference/object_loader.h:8-18) over the
// CUDA voxeliser: same name, arguments, defaults and console lines; the mesh is read on the host and
// everything else (rotation, rejection grid, ray-parity tests, obstacle mask) runs on the B200
// (include/fluid_cuda.h: fluid_voxelize).
#ifndef FLUID_B200_OBJECT_LOADER_H
#define FLUID_B200_OBJECT_LOADER_H

#include <string>
#include <vector>

class Simulation;

void loadSTLIntoObstacles(const char* stlFile, Simulation& sim, float scale = 0.8f, float rot_x = 0.0f,
                          float rot_y = 0.0f, float rot_z = 0.0f, float translate_x = 0.0f,
                          float translate_y = 0.0f, float translate_z = 0.0f);

/* the same with an explicit sampling rule (FLUID_VOX_HASHED / FLUID_VOX_FIXED / FLUID_VOX_CELLS); returns the
 * "Added N obstacle points" count, 0 when the file could not be read */
long long loadSTLIntoObstacles(const char* stlFile, Simulation& sim, int sampling, float scale, float rot_x,
                               float rot_y, float rot_z, float translate_x, float translate_y, float translate_z);

/* triangles of an ASCII or binary STL as 9 floats each (v1, v2, v3), object_loader.cpp:98-174 */
std::vector<float> readSTLTriangles(const std::string& filename);

#endif
