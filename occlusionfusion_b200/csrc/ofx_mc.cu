// ofx_mc.cu -- marching cubes on the device: the step the reference hands to scikit-image
// (`measure.marching_cubes(tsdf_vol, mask=diff_tsdf, level=0)`, fusion_with_occlusion/tsdf.py:794) plus the vertex colour
// unpack that follows it (tsdf.py:797-807).  Classic form: one case-table lookup per cube (csrc/ofx_mc_tables.h, derived
// by tools/gen_mc_tables.py), vertices on the grid edges at the linear zero crossing.  scikit-image is absent here, so this
// is checked against the CPU restatement of the same contract (the marching-cubes file of the test-only CPU checker), not against the reference's output.
//   Three streaming passes, no atomics, deterministic output order:
//     ofx_mc_count     per z-row (x*Dy + y): number of triangles (nothing per voxel is stored; 1 B of mask read per voxel,
//                      the 8 corners only where the mask holds)
//     [exclusive prefix sum over the Dx*Dy rows: caller, e.g. torch.cumsum]
//     ofx_mc_emit      per row: recomputes the cases, block scan along z, writes the triangles as grid-edge keys
//                      (3 * lower-corner index + axis) at row offset + in-row prefix
//     [unique keys = vertices, inverse = faces: caller, e.g. torch.unique]
//     ofx_mc_vertices  per unique key: position (voxel coordinates), gradient normal, unpacked colour
// A block walks z-rows (row = x*Dy + y: one division per row), threads run along z (coalesced).
#include "../../include/ofx.h"
#include "ofx_common.h"
#include "ofx_mc_tables.h"

namespace ofx {

int sm_count();   // ofx_points.cu

__device__ __forceinline__ int mc_case(const float *__restrict__ vol, const uint8_t *__restrict__ mask, int64_t i, int x, int y,
                                       int z, int W, int H, int D, float level)
{
    if (x + 1 >= W || y + 1 >= H || z + 1 >= D) return 0;
    if (mask && !__ldg(mask + i)) return 0;
    const int64_t sx = (int64_t)H * D, sy = D;
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

// pass 1: triangles per z-row (row = x*Dy + y).  Nothing per voxel is stored: pass 2 recomputes the cases.
__global__ void __launch_bounds__(256) mc_count_kernel(const float *__restrict__ vol, const uint8_t *__restrict__ mask, int W, int H,
                                                       int D, float level, int32_t *__restrict__ row_count)
{
    __shared__ int s_warp[8];
    for (int r = blockIdx.x; r < W * H; r += gridDim.x) {
        const int x = r / H, y = r - x * H;
        int mine = 0;
        for (int z = threadIdx.x; z < D; z += blockDim.x)
            mine += OFX_MC_TRIS[mc_case(vol, mask, (int64_t)r * D + z, x, y, z, W, H, D, level)];
        int total;
        block_exclusive_scan(mine, s_warp, &total);
        if (threadIdx.x == 0) row_count[r] = total;
    }
}

// pass 2: a block per z-row writes the row's triangles at the row's offset, cubes in z order (block scan per 256-cube chunk)
__global__ void __launch_bounds__(256) mc_emit_kernel(const float *__restrict__ vol, const uint8_t *__restrict__ mask, int W, int H,
                                                      int D, float level, const int32_t *__restrict__ row_count,
                                                      const int32_t *__restrict__ row_offset, int64_t *__restrict__ keys)
{
    __shared__ int s_warp[8];
    for (int r = blockIdx.x; r < W * H; r += gridDim.x) {
        if (__ldg(row_count + r) == 0) continue;                      // uniform across the block
        const int x = r / H, y = r - x * H;
        int carry = __ldg(row_offset + r);
        for (int z0 = 0; z0 < D; z0 += blockDim.x) {
            const int z = z0 + threadIdx.x;
            const int64_t i = (int64_t)r * D + z;
            const int c = z < D ? mc_case(vol, mask, i, x, y, z, W, H, D, level) : 0;
            const int nt = OFX_MC_TRIS[c];
            int total;
            const int before = block_exclusive_scan(nt, s_warp, &total);
            int64_t *out = keys + 3 * (int64_t)(carry + before);
            for (int k = 0; k < 3 * nt; ++k) {
                const int e = OFX_MC_EDGES[c][k], axis = e >> 2, b = e & 3;
                // lower corner of edge e: offsets (b & 1, b >> 1) along the two other axes, in increasing axis order
                const int ox = axis == 0 ? 0 : (b & 1), oz = axis == 2 ? 0 : (b >> 1);
                const int oy = axis == 0 ? (b & 1) : (axis == 1 ? 0 : (b >> 1));
                out[k] = 3 * (i + ((int64_t)ox * H + oy) * D + oz) + axis;
            }
            carry += total;
        }
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

__global__ void __launch_bounds__(256) mc_vertices_kernel(const float *__restrict__ vol, const float *__restrict__ colour, int W, int H,
                                                          int D, float level, const int64_t *__restrict__ keys, int64_t n,
                                                          float *__restrict__ verts, float *__restrict__ normals,
                                                          uint8_t *__restrict__ rgb)
{
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        const int64_t key = __ldg(keys + j), lower = key / 3;
        const int axis = (int)(key - 3 * lower);
        const int z = (int)(lower % D), y = (int)((lower / D) % H), x = (int)(lower / ((int64_t)D * H));
        const int64_t stride = axis == 0 ? (int64_t)H * D : (axis == 1 ? D : 1);
        const float v0 = __ldg(vol + lower), v1 = __ldg(vol + lower + stride);
        const float t = (level - v0) / (v1 - v0);
        float p[3] = {(float)x, (float)y, (float)z};
        p[axis] += t;
        verts[3 * j] = p[0]; verts[3 * j + 1] = p[1]; verts[3 * j + 2] = p[2];
        if (normals) {
            float g0[3], g1[3];
            mc_gradient(vol, x, y, z, W, H, D, g0);
            mc_gradient(vol, x + (axis == 0), y + (axis == 1), z + (axis == 2), W, H, D, g1);
            float g[3], s = 0.0f;
#pragma unroll
            for (int a = 0; a < 3; ++a) { g[a] = g0[a] + t * (g1[a] - g0[a]); s += g[a] * g[a]; }
            const float inv = s > 0.0f ? 1.0f / sqrtf(s) : 0.0f;       // towards increasing values: inside -> outside
            normals[3 * j] = g[0] * inv; normals[3 * j + 1] = g[1] * inv; normals[3 * j + 2] = g[2] * inv;
        }
        if (rgb) {
            // tsdf.py:797-807: colour of the voxel nearest to the vertex (np.round: half to even), unpacked b, g, r
            const int ix = (int)rintf(p[0]), iy = (int)rintf(p[1]), iz = (int)rintf(p[2]);
            const float c = __ldg(colour + ((int64_t)ix * H + iy) * D + iz);
            const float b = floorf(c / 65536.0f);
            const float g = floorf((c - b * 65536.0f) / 256.0f);
            const float r = c - b * 65536.0f - g * 256.0f;
            rgb[3 * j] = (uint8_t)floorf(r); rgb[3 * j + 1] = (uint8_t)floorf(g); rgb[3 * j + 2] = (uint8_t)floorf(b);
        }
    }
}

}  // namespace ofx

using namespace ofx;

static int mc_check(const float *vol, int32_t Dx, int32_t Dy, int32_t Dz, const char *who)
{
    OFX_REQUIRE(Dx >= 0 && Dy >= 0 && Dz >= 0 && (int64_t)Dx * Dy < 2147483647LL, "%s: bad dimensions", who);
    OFX_REQUIRE((int64_t)Dx * Dy * Dz == 0 || vol, "%s: null volume", who);
    return 0;
}

extern "C" int ofx_mc_count(const float *vol, const uint8_t *mask, int32_t Dx, int32_t Dy, int32_t Dz, float level,
                            int32_t *row_count, void *stream)
{
    if (mc_check(vol, Dx, Dy, Dz, "ofx_mc_count")) return 1;
    if ((int64_t)Dx * Dy * Dz == 0) return 0;
    OFX_REQUIRE(row_count, "ofx_mc_count: null pointer");
    const int rows = Dx * Dy, grid = rows < sm_count() * 8 ? rows : sm_count() * 8;
    mc_count_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(vol, mask, Dx, Dy, Dz, level, row_count);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_mc_emit(const float *vol, const uint8_t *mask, int32_t Dx, int32_t Dy, int32_t Dz, float level,
                           const int32_t *row_count, const int32_t *row_offset, int64_t *keys, void *stream)
{
    if (mc_check(vol, Dx, Dy, Dz, "ofx_mc_emit")) return 1;
    if ((int64_t)Dx * Dy * Dz == 0) return 0;
    OFX_REQUIRE(row_count && row_offset && keys, "ofx_mc_emit: null pointer");
    const int rows = Dx * Dy, grid = rows < sm_count() * 8 ? rows : sm_count() * 8;
    mc_emit_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(vol, mask, Dx, Dy, Dz, level, row_count, row_offset, keys);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_mc_vertices(const float *vol, const float *colour, int32_t Dx, int32_t Dy, int32_t Dz, float level,
                               const int64_t *keys, int64_t n, float *verts, float *normals, uint8_t *rgb, void *stream)
{
    if (mc_check(vol, Dx, Dy, Dz, "ofx_mc_vertices")) return 1;
    if (n == 0) return 0;
    OFX_REQUIRE(keys && verts && n > 0 && (!rgb || colour), "ofx_mc_vertices: bad arguments");
    int64_t blocks = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
    mc_vertices_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(vol, colour, Dx, Dy, Dz, level, keys, n,
                                                                                            verts, normals, rgb);
    OFX_LAUNCH_CHECK();
    return 0;
}
