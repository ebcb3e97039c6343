// ofx_mc.cu -- marching cubes on the device: the step the reference hands to scikit-image
// (`measure.marching_cubes(tsdf_vol, mask=diff_tsdf, level=0)`, fusion_with_occlusion/tsdf.py:794) plus the vertex colour
// unpack that follows it (tsdf.py:797-807).  Classic form: one case-table lookup per cube (csrc/ofx_mc_tables.h, derived
// by tools/gen_mc_tables.py), vertices on the grid edges at the linear zero crossing.  scikit-image is absent here, so this
// is checked against the CPU restatement of the same contract (the marching-cubes file of the test-only CPU checker), not against the reference's output.
//
// One library call (ofx_mc_mesh), four streaming passes of own kernels + two device scans, no sort, no atomics, no host
// read in between, deterministic output: vertices ascending by grid edge (3 * lower-corner index + axis), faces in cube order.
//   The sort of the earlier version (triangle corners -> unique grid edges) is replaced by OWNERSHIP: a grid edge has a
//   vertex exactly when its end points lie on different sides of the level and one of the (up to four) cubes around it is
//   processed -- the case table uses every crossed edge of a cube (tests/test_mc_oracle.py checks the table for that).  So
//   the vertices of a z-row (row = x*Dy + y) can be counted and numbered from the volume alone, in key order:
//     mc_row_flags_kernel per row: does the mask hold a cube at all (the truncated region is sparse: the count pass then
//                         reads the volume only around flagged rows)
//     mc_count_kernel     per row: triangles of its cubes, vertices of the edges whose lower corner lies in the row
//     exclusive_scan x 2  row offsets of triangles and vertices; the totals stay on the device (counts[2])
//     mc_vertices_kernel  per row: position, gradient normal, colour of every owned vertex at row offset + rank, and its
//                         code z*3 + axis (u16) beside it
//     mc_faces_kernel     per row: each triangle corner = offset of the owning row + rank of the corner's code among that
//                         row's codes (binary search over a few dozen entries)
// A block walks z-rows (one division per row), threads run along z (coalesced).
#include "../../include/ofx.h"
#include "ofx_common.h"
#include "ofx_mc_tables.h"
#include "ofx_scan.cuh"

namespace ofx {

int sm_count();                                                   // ofx_points.cu
int graph_scratch_alloc(void **p, size_t bytes, cudaStream_t s);  // ofx_graph.cu: the stream-ordered scratch pool

struct McGrid {
    const float *vol;
    const uint8_t *mask;
    int W, H, D;
    float level;
};

__device__ __forceinline__ bool cube_processed(const McGrid &g, int x, int y, int z)
{
    if (x < 0 || y < 0 || z < 0 || x + 1 >= g.W || y + 1 >= g.H || z + 1 >= g.D) return false;
    return !g.mask || __ldg(g.mask + ((int64_t)x * g.H + y) * g.D + z) != 0;
}

__device__ __forceinline__ int mc_case(const McGrid &g, int64_t i, int x, int y, int z)
{
    if (!cube_processed(g, x, y, z)) return 0;
    const int64_t sx = (int64_t)g.H * g.D, sy = g.D;
    const float *vol = g.vol;
    const float level = g.level;
    int c = 0;
    c |= (__ldg(vol + i) < level) << 0;
    c |= (__ldg(vol + i + sx) < level) << 1;
    c |= (__ldg(vol + i + sy) < level) << 2;
    c |= (__ldg(vol + i + sx + sy) < level) << 3;
    c |= (__ldg(vol + i + 1) < level) << 4;
    c |= (__ldg(vol + i + sx + 1) < level) << 5;
    c |= (__ldg(vol + i + sy + 1) < level) << 6;
    c |= (__ldg(vol + i + sx + sy + 1) < level) << 7;
    return c;
}

// bit a set: the grid edge from (x,y,z) along axis a carries a vertex (crossed, and a processed cube touches it)
__device__ __forceinline__ int owned_edges(const McGrid &g, int64_t i, int x, int y, int z)
{
    const int64_t sx = (int64_t)g.H * g.D, sy = g.D;
    const bool in0 = __ldg(g.vol + i) < g.level;
    int m = 0;
    if (x + 1 < g.W && (__ldg(g.vol + i + sx) < g.level) != in0)
        if (cube_processed(g, x, y, z) || cube_processed(g, x, y - 1, z) || cube_processed(g, x, y, z - 1) || cube_processed(g, x, y - 1, z - 1))
            m |= 1;
    if (y + 1 < g.H && (__ldg(g.vol + i + sy) < g.level) != in0)
        if (cube_processed(g, x, y, z) || cube_processed(g, x - 1, y, z) || cube_processed(g, x, y, z - 1) || cube_processed(g, x - 1, y, z - 1))
            m |= 2;
    if (z + 1 < g.D && (__ldg(g.vol + i + 1) < g.level) != in0)
        if (cube_processed(g, x, y, z) || cube_processed(g, x - 1, y, z) || cube_processed(g, x, y - 1, z) || cube_processed(g, x - 1, y - 1, z))
            m |= 4;
    return m;
}

// block-wide exclusive scan of one int per thread (256 threads); returns the exclusive prefix, *total = block sum
__device__ __forceinline__ int block_exclusive_scan(int v, int *s_warp, int *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    __syncthreads();                       // s_warp may still be read by the previous call
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int before = 0, sum = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { const int t = s_warp[w]; if (w < warp) before += t; sum += t; }
    *total = sum;
    return before + incl - v;
}

// pass 0 (only with a mask): row_flag[r] = the z-row r holds a processed cube.  The mask is sparse (the truncated region
// is a fraction of a percent of a fusion volume): the count pass then touches the volume only around flagged rows.
__global__ void __launch_bounds__(256) mc_row_flags_kernel(const uint8_t *__restrict__ mask, int rows, int D, uint8_t *__restrict__ row_flag)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    for (int r = warp; r < rows; r += n_warps) {
        const uint8_t *m = mask + (int64_t)r * D;
        int any = 0;
        if ((D & 15) == 0 && (reinterpret_cast<uintptr_t>(m) & 15) == 0) {
            for (int q = lane; q < (D >> 4); q += 32) {
                const uint4 w = __ldg(reinterpret_cast<const uint4 *>(m) + q);
                any |= (w.x | w.y | w.z | w.w) != 0u;
            }
        } else {
            for (int z = lane; z < D; z += 32) any |= __ldg(m + z) != 0;
        }
        any = __any_sync(0xffffffffu, any);
        if (lane == 0) row_flag[r] = (uint8_t)any;
    }
}

// pass 1: triangles and owned vertices per z-row.  Nothing per voxel is stored: the later passes recompute.
// row_flag == nullptr (no mask): every row counts.
__global__ void __launch_bounds__(256) mc_count_kernel(McGrid g, const uint8_t *__restrict__ row_flag, int32_t *__restrict__ row_tris,
                                                       int32_t *__restrict__ row_verts)
{
    __shared__ int s_warp[8];
    for (int r = blockIdx.x; r < g.W * g.H; r += gridDim.x) {
        const int x = r / g.H, y = r - x * g.H;
        // cubes of this row make triangles; cubes of the rows (x-1..x, y-1..y) touch the edges that start in this row
        bool f_t = true, f_v = true;
        if (row_flag) {
            f_t = row_flag[r] != 0;
            f_v = f_t || (x > 0 && row_flag[r - g.H]) || (y > 0 && row_flag[r - 1]) || (x > 0 && y > 0 && row_flag[r - g.H - 1]);
        }
        if (!f_v) {                                                    // uniform across the block
            if (threadIdx.x == 0) { row_tris[r] = 0; row_verts[r] = 0; }
            continue;
        }
        int tris = 0, verts = 0;
        for (int z = threadIdx.x; z < g.D; z += blockDim.x) {
            const int64_t i = (int64_t)r * g.D + z;
            if (f_t) tris += OFX_MC_TRIS[mc_case(g, i, x, y, z)];
            verts += __popc(owned_edges(g, i, x, y, z));
        }
        int total_t, total_v;
        block_exclusive_scan(tris, s_warp, &total_t);
        block_exclusive_scan(verts, s_warp, &total_v);
        if (threadIdx.x == 0) { row_tris[r] = total_t; row_verts[r] = total_v; }
    }
}

// central-difference gradient of the volume at a grid point (one-sided at the volume boundary)
__device__ __forceinline__ void mc_gradient(const float *__restrict__ vol, int x, int y, int z, int W, int H, int D, float g[3])
{
    const int64_t sx = (int64_t)H * D, sy = D, i = ((int64_t)x * H + y) * D + z;
    g[0] = __ldg(vol + i + (x + 1 < W ? sx : 0)) - __ldg(vol + i - (x > 0 ? sx : 0));
    g[1] = __ldg(vol + i + (y + 1 < H ? sy : 0)) - __ldg(vol + i - (y > 0 ? sy : 0));
    g[2] = __ldg(vol + i + (z + 1 < D ? 1 : 0)) - __ldg(vol + i - (z > 0 ? 1 : 0));
}

__device__ __forceinline__ void mc_write_vertex(const McGrid &g, const float *__restrict__ colour, int x, int y, int z, int axis,
                                                int64_t j, float *__restrict__ verts, float *__restrict__ normals,
                                                uint8_t *__restrict__ rgb)
{
    const int W = g.W, H = g.H, D = g.D;
    const int64_t lower = ((int64_t)x * H + y) * D + z;
    const int64_t stride = axis == 0 ? (int64_t)H * D : (axis == 1 ? D : 1);
    const float v0 = __ldg(g.vol + lower), v1 = __ldg(g.vol + lower + stride);
    const float t = (g.level - v0) / (v1 - v0);
    float p[3] = {(float)x, (float)y, (float)z};
    p[axis] += t;
    verts[3 * j] = p[0]; verts[3 * j + 1] = p[1]; verts[3 * j + 2] = p[2];
    if (normals) {
        float g0[3], g1[3];
        mc_gradient(g.vol, x, y, z, W, H, D, g0);
        mc_gradient(g.vol, x + (axis == 0), y + (axis == 1), z + (axis == 2), W, H, D, g1);
        float gr[3], s = 0.0f;
#pragma unroll
        for (int a = 0; a < 3; ++a) { gr[a] = g0[a] + t * (g1[a] - g0[a]); s += gr[a] * gr[a]; }
        const float inv = s > 0.0f ? 1.0f / sqrtf(s) : 0.0f;       // towards increasing values: inside -> outside
        normals[3 * j] = gr[0] * inv; normals[3 * j + 1] = gr[1] * inv; normals[3 * j + 2] = gr[2] * inv;
    }
    if (rgb) {
        // tsdf.py:797-807: colour of the voxel nearest to the vertex (np.round: half to even), unpacked b, g, r
        const int ix = (int)rintf(p[0]), iy = (int)rintf(p[1]), iz = (int)rintf(p[2]);
        const float c = __ldg(colour + ((int64_t)ix * H + iy) * D + iz);
        const float b = floorf(c / 65536.0f);
        const float gch = floorf((c - b * 65536.0f) / 256.0f);
        const float rch = c - b * 65536.0f - gch * 256.0f;
        rgb[3 * j] = (uint8_t)floorf(rch); rgb[3 * j + 1] = (uint8_t)floorf(gch); rgb[3 * j + 2] = (uint8_t)floorf(b);
    }
}

// pass 2: the owned vertices of every row, in key order, at the row's offset; vcode[j] = z*3 + axis of vertex j
__global__ void __launch_bounds__(256) mc_vertices_kernel(McGrid g, const float *__restrict__ colour,
                                                          const int32_t *__restrict__ row_verts, const int32_t *__restrict__ vert_off,
                                                          const int32_t *__restrict__ counts, int32_t cap_verts, int32_t cap_tris,
                                                          float *__restrict__ verts, float *__restrict__ normals,
                                                          uint8_t *__restrict__ rgb, uint16_t *__restrict__ vcode)
{
    __shared__ int s_warp[8];
    if (counts[0] > cap_verts || counts[1] > cap_tris) return;          // the caller reads counts and comes back with room
    for (int r = blockIdx.x; r < g.W * g.H; r += gridDim.x) {
        if (__ldg(row_verts + r) == 0) continue;                        // uniform across the block
        const int x = r / g.H, y = r - x * g.H;
        int carry = __ldg(vert_off + r);
        for (int z0 = 0; z0 < g.D; z0 += blockDim.x) {
            const int z = z0 + threadIdx.x;
            const int m = z < g.D ? owned_edges(g, (int64_t)r * g.D + z, x, y, z) : 0;
            int total;
            int64_t j = carry + block_exclusive_scan(__popc(m), s_warp, &total);
#pragma unroll
            for (int a = 0; a < 3; ++a)
                if (m >> a & 1) {
                    mc_write_vertex(g, colour, x, y, z, a, j, verts, normals, rgb);
                    vcode[j] = (uint16_t)(z * 3 + a);
                    ++j;
                }
            carry += total;
        }
    }
}

// pass 3: the triangles of every row at the row's offset, cubes in z order; corner -> vertex index
__global__ void __launch_bounds__(256) mc_faces_kernel(McGrid g, const int32_t *__restrict__ row_tris, const int32_t *__restrict__ tri_off,
                                                       const int32_t *__restrict__ vert_off, const uint16_t *__restrict__ vcode,
                                                       const int32_t *__restrict__ counts, int32_t cap_verts, int32_t cap_tris,
                                                       int32_t *__restrict__ faces)
{
    __shared__ int s_warp[8];
    if (counts[0] > cap_verts || counts[1] > cap_tris) return;
    for (int r = blockIdx.x; r < g.W * g.H; r += gridDim.x) {
        if (__ldg(row_tris + r) == 0) continue;                         // uniform across the block
        const int x = r / g.H, y = r - x * g.H;
        int carry = __ldg(tri_off + r);
        for (int z0 = 0; z0 < g.D; z0 += blockDim.x) {
            const int z = z0 + threadIdx.x;
            const int c = z < g.D ? mc_case(g, (int64_t)r * g.D + z, x, y, z) : 0;
            const int nt = OFX_MC_TRIS[c];
            int total;
            const int before = block_exclusive_scan(nt, s_warp, &total);
            int32_t *out = faces + 3 * (int64_t)(carry + before);
            for (int k = 0; k < 3 * nt; ++k) {
                const int e = OFX_MC_EDGES[c][k], axis = e >> 2, b = e & 3;
                // lower corner of edge e: offsets (b & 1, b >> 1) along the two other axes, in increasing axis order
                const int ox = axis == 0 ? 0 : (b & 1), oz = axis == 2 ? 0 : (b >> 1);
                const int oy = axis == 0 ? (b & 1) : (axis == 1 ? 0 : (b >> 1));
                const int ro = r + ox * g.H + oy;                       // owning row
                const int code = (z + oz) * 3 + axis;
                int lo = __ldg(vert_off + ro), hi = __ldg(vert_off + ro + 1);        // vert_off has rows + 1 entries
                while (lo < hi) {                                                     // first entry >= code (it is there)
                    const int mid = (lo + hi) >> 1;
                    if ((int)__ldg(vcode + mid) < code) lo = mid + 1; else hi = mid;
                }
                out[k] = lo;
            }
            carry += total;
        }
    }
}

}  // namespace ofx

using namespace ofx;

extern "C" int ofx_mc_mesh(const float *vol, const uint8_t *mask, const float *colour, int32_t Dx, int32_t Dy, int32_t Dz,
                           float level, int32_t cap_verts, int32_t cap_tris, float *verts, float *normals, uint8_t *rgb,
                           int32_t *faces, int32_t *counts, void *stream)
{
    OFX_REQUIRE(Dx >= 0 && Dy >= 0 && Dz >= 0 && (int64_t)Dx * Dy < 2147483647LL, "ofx_mc_mesh: bad dimensions");
    OFX_REQUIRE(Dz <= 21845, "ofx_mc_mesh: Dz = %d exceeds 21845 (vertex codes are 16 bits)", Dz);
    OFX_REQUIRE(counts && cap_verts >= 0 && cap_tris >= 0, "ofx_mc_mesh: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    OFX_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(int32_t), s));
    if ((int64_t)Dx * Dy * Dz == 0) return 0;
    OFX_REQUIRE(vol && (!rgb || colour), "ofx_mc_mesh: null volume");
    OFX_REQUIRE((cap_verts == 0 || verts) && (cap_tris == 0 || faces), "ofx_mc_mesh: null output");
    const int rows = Dx * Dy, grid = rows < sm_count() * 8 ? rows : sm_count() * 8;
    const int n_sums = (rows + SCAN_BLOCK * SCAN_ITEMS - 1) / (SCAN_BLOCK * SCAN_ITEMS) + 2;
    // scratch (stream-ordered pool): four row arrays, the scan's block sums, one 16-bit code per vertex
    int32_t *rowbuf = nullptr;
    uint16_t *vcode = nullptr;
    const size_t n_row_ints = (size_t)4 * rows + 2 + n_sums + (mask ? (rows + 3) / 4 + 1 : 0);      // + one flag byte per row
    if (graph_scratch_alloc((void **)&rowbuf, n_row_ints * sizeof(int32_t), s)) return 1;
    if (graph_scratch_alloc((void **)&vcode, ((size_t)cap_verts + 8) * sizeof(uint16_t), s)) { cudaFreeAsync(rowbuf, s); return 1; }
    int32_t *row_tris = rowbuf, *row_verts = rowbuf + rows, *tri_off = rowbuf + 2 * rows, *vert_off = rowbuf + 3 * rows + 1;
    int32_t *sums = rowbuf + 4 * rows + 2;                                // tri_off and vert_off have rows + 1 entries
    uint8_t *row_flag = mask ? reinterpret_cast<uint8_t *>(rowbuf + 4 * rows + 2 + n_sums) : nullptr;
    const McGrid g{vol, mask, Dx, Dy, Dz, level};
    int rc = 0;
    if (mask) mc_row_flags_kernel<<<sm_count() * 8, 256, 0, s>>>(mask, rows, Dz, row_flag);
    mc_count_kernel<<<grid, 256, 0, s>>>(g, row_flag, row_tris, row_verts);
    if (!rc) rc = exclusive_scan(row_verts, rows, vert_off, sums, counts + 0, s);
    if (!rc) rc = exclusive_scan(row_tris, rows, tri_off, sums, counts + 1, s);
    if (!rc && cudaMemcpyAsync(vert_off + rows, counts, sizeof(int32_t), cudaMemcpyDeviceToDevice, s) != cudaSuccess) rc = 1;
    if (!rc) {
        mc_vertices_kernel<<<grid, 256, 0, s>>>(g, colour, row_verts, vert_off, counts, cap_verts, cap_tris, verts, normals, rgb, vcode);
        mc_faces_kernel<<<grid, 256, 0, s>>>(g, row_tris, tri_off, vert_off, vcode, counts, cap_verts, cap_tris, faces);
    }
    cudaFreeAsync(rowbuf, s);
    cudaFreeAsync(vcode, s);
    if (rc) return rc;
    OFX_LAUNCH_CHECK();
    return 0;
}
