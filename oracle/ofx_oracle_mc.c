/* ofx_oracle_mc.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the marching-cubes step of the reference's surface extraction (fusion_with_occlusion/tsdf.py:794:
 * `measure.marching_cubes(tsdf_vol, mask=diff_tsdf, level=0)`).  PARITY UNPINNED: the reference calls scikit-image's
 * Lewiner variant; scikit-image is not installed here and not vendored in /root/reference, so there is no output of the
 * reference to pin against.  What is restated is the published algorithm in its classic form (one case table lookup per
 * cube, vertices on the grid edges at the linear zero crossing), with the case table derived by tools/gen_mc_tables.py
 * (face-consistent: watertight, ambiguous faces always separate the inside corners).  Lewiner's variant adds interior
 * tests for the ambiguous cases, so triangulations can differ there; vertex positions on edges are the same formula.
 *
 * Contract shared with csrc/ofx_mc.cu:
 *   cubes (x,y,z), x < W-1, y < H-1, z < D-1, in linear-index order i = (x*H + y)*D + z; a cube is processed when
 *   mask == NULL or mask[i] != 0 (its lowest corner); corner c = x + 2y + 4z offsets, inside <=> value < level;
 *   a triangle corner is named by its grid edge: key = 3 * (linear index of the edge's lower corner) + axis.
 * This file emits the triangle keys; oracle.py sorts them into unique vertices and places the vertices.
 */
#include <stdint.h>
#include "../occlusionfusion_b200/csrc/ofx_mc_tables.h"

/* keys == NULL: count only.  Returns the number of triangles. */
int64_t ofx_oracle_mc_triangles(const float *vol, const uint8_t *mask, int W, int H, int D, float level, int64_t *keys)
{
    int64_t nt = 0;
    for (int x = 0; x + 1 < W; ++x)
        for (int y = 0; y + 1 < H; ++y)
            for (int z = 0; z + 1 < D; ++z) {
                const int64_t i = ((int64_t)x * H + y) * D + z;
                if (mask && !mask[i]) continue;
                int c = 0;
                for (int k = 0; k < 8; ++k) {
                    const int64_t j = i + ((int64_t)(k & 1) * H + ((k >> 1) & 1)) * D + (k >> 2);
                    if (vol[j] < level) c |= 1 << k;
                }
                for (int t = 0; t < OFX_MC_TRIS[c]; ++t, ++nt) {
                    if (!keys) continue;
                    for (int v = 0; v < 3; ++v) {
                        const int e = OFX_MC_EDGES[c][3 * t + v], axis = e >> 2, r = e & 3;
                        /* lower corner of edge e: offsets (b0, b1) along the two other axes, in increasing axis order */
                        int off[3] = {0, 0, 0};
                        off[axis == 0 ? 1 : 0] = r & 1;
                        off[axis == 2 ? 1 : 2] = r >> 1;
                        const int64_t lower = i + ((int64_t)off[0] * H + off[1]) * D + off[2];
                        keys[3 * nt + v] = 3 * lower + axis;
                    }
                }
            }
    return nt;
}
