/* oracle/mesh_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of what GUI/utils.py:10-38 (generate_obstacle_mesh) asks for:
 *     verts, faces, _, _ = measure.marching_cubes(obs_data, level=0.5)          (utils.py:17)
 * on the padded obstacle volume in [x, y, z] index order (GUI/main_window.py:204).  The checker for
 * csrc/mesh.cuh.
 *
 * PARITY UNPINNED.  The algorithm lives in a third-party dependency, scikit-image (`from skimage import
 * measure`, utils.py:6; the reference pins no version -- it has no requirements file), which is absent from
 * /root/reference and from this image.  Its marching_cubes defaults to the Lewiner variant (Lewiner et al.
 * 2003, "Efficient implementation of marching cubes' cases with topological guarantees"): the classic
 * per-cube table extended by face and interior tests that choose between triangulations of the ambiguous
 * cases, vertices placed by linear interpolation along the crossed grid edges, faces oriented so that normals
 * point towards lower values (gradient_direction='descent').  What is restated here is that published
 * scheme with a PLAIN case table (fluid_simulation_b200/csrc/mc_table.h, generated from first principles by
 * scripts/make_mc_table.py): it yields the same vertex set (one vertex per crossed edge; Lewiner adds an
 * interior vertex only in some ambiguous configurations) and a closed surface with the same orientation, but
 * not the same triangulation of ambiguous cubes, and not skimage's vertex order.  There is no golden output
 * of the reference's call to compare with; tests/test_mesh.py checks the properties the data determines.
 *
 * Serial and deliberately unlike the device code: a first sweep numbers the crossed edges in a hash-free
 * "edge id" volume (three ints per cell), a second sweep walks the cubes. */
#include <stdlib.h>
#include <string.h>

#define MC_TABLE_QUAL static const
#include "../fluid_simulation_b200/csrc/mc_table.h"

typedef struct { const float* a; int nx, ny, nz; } mvol_t;

static float mat(const mvol_t* g, int x, int y, int z)
{
    return g->a[(size_t)x + (size_t)y * g->nx + (size_t)z * g->nx * g->ny];
}

/* obs: frame-file layout (x fastest).  Counts are always returned; verts (3 floats each) and faces (3 ints each)
 * are written when they are not NULL (the caller sizes them from a first call).  Returns 0, -1 on allocation
 * failure. */
int mo_marching_cubes(const float* obs, int nx, int ny, int nz, float level, long long* n_verts, long long* n_faces,
                      float* verts, int* faces)
{
    const mvol_t g = { obs, nx, ny, nz };
    const size_t ncell = (size_t)nx * ny * nz;
    int* edge_id = (int*)malloc(ncell * 3 * sizeof(int));
    if (!edge_id) return -1;
    long long nv = 0, nf = 0;
    static const int step[3][3] = { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
    for (int z = 0; z < nz; ++z)
        for (int y = 0; y < ny; ++y)
            for (int x = 0; x < nx; ++x) {
                const size_t c = (size_t)x + (size_t)y * nx + (size_t)z * nx * ny;
                const float v0 = mat(&g, x, y, z);
                for (int a = 0; a < 3; ++a) {
                    edge_id[3 * c + a] = -1;
                    const int x1 = x + step[a][0], y1 = y + step[a][1], z1 = z + step[a][2];
                    if (x1 >= nx || y1 >= ny || z1 >= nz) continue;
                    const float v1 = mat(&g, x1, y1, z1);
                    if ((v0 > level) == (v1 > level)) continue;
                    if (verts) {
                        const float t = (level - v0) / (v1 - v0);
                        verts[3 * nv + 0] = (float)x + (a == 0 ? t : 0.0f);
                        verts[3 * nv + 1] = (float)y + (a == 1 ? t : 0.0f);
                        verts[3 * nv + 2] = (float)z + (a == 2 ? t : 0.0f);
                    }
                    edge_id[3 * c + a] = (int)nv++;
                }
            }
    for (int z = 0; z + 1 < nz; ++z)
        for (int y = 0; y + 1 < ny; ++y)
            for (int x = 0; x + 1 < nx; ++x) {
                int kase = 0;
                for (int i = 0; i < 8; ++i)
                    if (mat(&g, x + (i & 1), y + ((i >> 1) & 1), z + ((i >> 2) & 1)) > level) kase |= 1 << i;
                for (int k = 0; MC_TRI[kase][k] >= 0; k += 3) {
                    if (faces)
                        for (int j = 0; j < 3; ++j) {
                            const int e = MC_TRI[kase][k + j], lo = MC_EDGE_CORNER[e][0];
                            const size_t c = (size_t)(x + (lo & 1)) + (size_t)(y + ((lo >> 1) & 1)) * nx +
                                             (size_t)(z + ((lo >> 2) & 1)) * nx * ny;
                            faces[3 * nf + j] = edge_id[3 * c + (e >> 2)];
                        }
                    ++nf;
                }
            }
    free(edge_id);
    *n_verts = nv; *n_faces = nf;
    return 0;
}
