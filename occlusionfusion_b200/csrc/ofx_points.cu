// ofx_points.cu -- free-standing point-set operators of libofx.so (sm_100a):
//   ofx_pack_frame        TSDFVolume.update           tsdf.py:545-572
//   ofx_skin              WarpField.skin              warpfield.py:83-129
//   ofx_warp_points       WarpField.deform            warpfield.py:270-305 (deform_lbs :208-231)
//   ofx_check_visibility  TSDFVolume.check_visibility tsdf.py:599-612
// These serve the callers either side of the volume path (mesh vertices, graph nodes, frame
// upload).  They are HBM/latency-bound streaming kernels: one thread per point (or pixel),
// coalesced loads, node table staged in shared memory, grid sized in multiples of the SM count.
#include "../../include/ofx.h"
#include "ofx_common.h"
#include "ofx_math.cuh"

namespace ofx {

// SM count of the CURRENT device (asked every time: the runtime caches device attributes, and a process may
// hold handles on several devices)
int sm_count()
{
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
    return n;
}

static inline int grid_for(int64_t n, int block, int ctas_per_sm)
{
    int64_t need = (n + block - 1) / block;
    int64_t cap = (int64_t)sm_count() * ctas_per_sm;
    return (int)(need < 1 ? 1 : (need < cap ? need : cap));
}

// ---- pack_frame ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_frame_kernel(const float *__restrict__ im, int depth_channel, int64_t npix,
                                                         float *__restrict__ colour, float *__restrict__ depth)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const float r = 255.0f * __ldg(im + i);
        const float g = 255.0f * __ldg(im + npix + i);
        const float b = 255.0f * __ldg(im + 2 * npix + i);
        float s = b * 65536.0f;
        const float t = g * 256.0f;
        s = s + t;
        s = s + r;
        colour[i] = floorf(s);                    // tsdf.py:562
        depth[i] = __ldg(im + depth_channel * npix + i);      // tsdf.py:554  depth = im[-1]
    }
}

// ---- skin ------------------------------------------------------------------------------------
constexpr int SKIN_NODE_TILE = 2048;   // 24 KB of node positions per stage

template <int K>
__global__ void __launch_bounds__(256) skin_points_kernel(const float *__restrict__ pts, int64_t n,
                                                          const float *__restrict__ nodes, int M, float cutoff,
                                                          float denom, int32_t *__restrict__ anchors,
                                                          float *__restrict__ weights, uint8_t *__restrict__ valid)
{
    __shared__ float s_nodes[SKIN_NODE_TILE * 3];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t n_round = ((n + stride - 1) / stride) * stride;   // keep whole CTAs in the loop for __syncthreads
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool live = i < n;
        float px = 0.f, py = 0.f, pz = 0.f;
        if (live) { px = __ldg(pts + 3 * i); py = __ldg(pts + 3 * i + 1); pz = __ldg(pts + 3 * i + 2); }
        float bd[K];
        int bi[K];
#pragma unroll
        for (int k = 0; k < K; ++k) { bd[k] = __int_as_float(0x7f800000); bi[k] = -1; }
        for (int base = 0; base < M; base += SKIN_NODE_TILE) {
            const int cnt = min(SKIN_NODE_TILE, M - base);
            __syncthreads();
            for (int t = threadIdx.x; t < cnt * 3; t += blockDim.x) s_nodes[t] = __ldg(nodes + (int64_t)base * 3 + t);
            __syncthreads();
            if (live) {
                for (int j = 0; j < cnt; ++j) {
                    const float d2 = dist2(px, py, pz, s_nodes[3 * j], s_nodes[3 * j + 1], s_nodes[3 * j + 2]);
                    topk_insert<K>(d2, base + j, bd, bi);
                }
            }
        }
        if (live) {
            float w[K];
            bool in_range[K];
            const bool ok = skin_weights<K>(bd, cutoff, denom, w, in_range);
#pragma unroll
            for (int k = 0; k < K; ++k) {
                anchors[K * i + k] = in_range[k] ? bi[k] : -1;   // warpfield.py:113
                weights[K * i + k] = w[k];
            }
            valid[i] = ok ? 1 : 0;                               // warpfield.py:120
        }
    }
}

// ---- warp_points -------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) warp_points_kernel(const float *__restrict__ pts, int64_t n,
                                                          const int32_t *__restrict__ anchors,
                                                          const float *__restrict__ weights,
                                                          const uint8_t *__restrict__ valid, int K,
                                                          const float *__restrict__ R, const float *__restrict__ T,
                                                          float *__restrict__ out)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float px = __ldg(pts + 3 * i), py = __ldg(pts + 3 * i + 1), pz = __ldg(pts + 3 * i + 2);
        float ax = px, ay = py, az = pz;
        if (valid[i]) {
            ax = ay = az = 0.0f;
            for (int k = 0; k < K; ++k) {
                const float w = __ldg(weights + K * i + k);
                if (w == 0.0f) continue;                          // warpfield.py:219
                const int a = __ldg(anchors + K * i + k);
                float rec[12];
#pragma unroll
                for (int e = 0; e < 9; ++e) rec[e] = __ldg(R + 9 * (int64_t)a + e);
#pragma unroll
                for (int e = 0; e < 3; ++e) rec[9 + e] = __ldg(T + 3 * (int64_t)a + e);
                lbs_accumulate(rec, w, px, py, pz, ax, ay, az);
            }
        }
        out[3 * i] = ax; out[3 * i + 1] = ay; out[3 * i + 2] = az;
    }
}

// ---- check_visibility ----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) check_visibility_kernel(const float *__restrict__ pts, int64_t n,
                                                               const float *__restrict__ depth, Camera cam,
                                                               double trunc, uint8_t *__restrict__ valid,
                                                               double *__restrict__ diff_out)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float qx = __ldg(pts + 3 * i), qy = __ldg(pts + 3 * i + 1), qz = __ldg(pts + 3 * i + 2);
        int u, v;
        double d = 0.0;                                           // np.zeros, tsdf.py:594
        if (project(qx, qy, qz, cam, u, v)) d = (double)__ldg(depth + (int64_t)v * cam.W + u);
        const double diff = d - (double)qz;                       // tsdf.py:608
        if (diff_out) diff_out[i] = diff;
        if (valid) valid[i] = (d > 0.0 && diff >= -trunc) ? 1 : 0;   // tsdf.py:609
    }
}

}  // namespace ofx

using namespace ofx;

extern "C" int ofx_pack_frame(const float *im, int32_t C, int32_t H, int32_t W, float *colour_packed, float *depth,
                              void *stream)
{
    OFX_REQUIRE(im && colour_packed && depth && H > 0 && W > 0 && C >= 4, "ofx_pack_frame: bad arguments");
    const int64_t npix = (int64_t)H * W;
    pack_frame_kernel<<<grid_for(npix, 256, 8), 256, 0, (cudaStream_t)stream>>>(im, C - 1, npix, colour_packed, depth);
    OFX_LAUNCH_CHECK();
    return 0;
}

template <int K>
static int launch_skin(const float *points, int64_t n, const float *nodes, int32_t M, float cutoff, float denom,
                       int32_t *anchors, float *weights, uint8_t *valid, cudaStream_t s)
{
    skin_points_kernel<K><<<grid_for(n, 256, 4), 256, 0, s>>>(points, n, nodes, M, cutoff, denom, anchors, weights, valid);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_skin(const float *points, int64_t n, const float *nodes, int32_t M, double node_coverage,
                        int32_t K, int32_t *anchors, float *weights, uint8_t *valid, void *stream)
{
    OFX_REQUIRE(K >= 1 && K <= OFX_MAX_K && M >= K, "ofx_skin: need 1 <= K <= %d and K <= M (K=%d, M=%d)", OFX_MAX_K, K, M);
    if (n == 0) return 0;
    OFX_REQUIRE(points && nodes && anchors && weights && valid, "ofx_skin: null pointer");
    const float cutoff = (float)(4.0 * node_coverage);                       // warpfield.py:109
    const float denom = (float)(2.0 * (node_coverage * node_coverage));      // warpfield.py:114
    cudaStream_t s = (cudaStream_t)stream;
    switch (K) {
    case 1: return launch_skin<1>(points, n, nodes, M, cutoff, denom, anchors, weights, valid, s);
    case 2: return launch_skin<2>(points, n, nodes, M, cutoff, denom, anchors, weights, valid, s);
    case 3: return launch_skin<3>(points, n, nodes, M, cutoff, denom, anchors, weights, valid, s);
    case 4: return launch_skin<4>(points, n, nodes, M, cutoff, denom, anchors, weights, valid, s);
    case 5: return launch_skin<5>(points, n, nodes, M, cutoff, denom, anchors, weights, valid, s);
    case 6: return launch_skin<6>(points, n, nodes, M, cutoff, denom, anchors, weights, valid, s);
    case 7: return launch_skin<7>(points, n, nodes, M, cutoff, denom, anchors, weights, valid, s);
    default: return launch_skin<8>(points, n, nodes, M, cutoff, denom, anchors, weights, valid, s);
    }
}

extern "C" int ofx_warp_points(const float *points, int64_t n, const int32_t *anchors, const float *weights,
                               const uint8_t *valid, int32_t K, const float *rotations, const float *translations,
                               float *out, void *stream)
{
    OFX_REQUIRE(K >= 1 && K <= OFX_MAX_K, "ofx_warp_points: K=%d out of range", K);
    if (n == 0) return 0;
    OFX_REQUIRE(points && anchors && weights && valid && rotations && translations && out, "ofx_warp_points: null pointer");
    warp_points_kernel<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(points, n, anchors, weights, valid, K,
                                                                             rotations, translations, out);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_check_visibility(const float *points, int64_t n, const float *depth, int32_t H, int32_t W,
                                    const float *intr_host, double trunc_margin, uint8_t *valid, double *depth_diff,
                                    void *stream)
{
    if (n == 0) return 0;
    OFX_REQUIRE(points && depth && intr_host && H > 0 && W > 0, "ofx_check_visibility: bad arguments");
    Camera cam{intr_host[0], intr_host[1], intr_host[2], intr_host[3], H, W};
    check_visibility_kernel<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(points, n, depth, cam, trunc_margin,
                                                                                  valid, depth_diff);
    OFX_LAUNCH_CHECK();
    return 0;
}

// ---- peer pull copy (multi-GPU frame packet) ---------------------------------------------------------
namespace ofx {
__global__ void __launch_bounds__(256) peer_copy_kernel(uint4 *__restrict__ dst, const uint4 *__restrict__ src, int64_t n16)
{
    // four independent 16-byte loads in flight per thread: NVLink round trips are ~2 us
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n16; i += 4 * stride) {
        const uint4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
        dst[i] = a; dst[i + stride] = b; dst[i + 2 * stride] = c; dst[i + 3 * stride] = d;
    }
    for (; i < n16; i += stride) dst[i] = src[i];
}
// One store per 16 bytes to a MULTICAST address (the NVSwitch replicates it into every mapped GPU's buffer): the frame packet
// leaves the sending GPU once instead of once per peer.
__global__ void __launch_bounds__(256) multicast_copy_kernel(float4 *__restrict__ mc_dst, const float4 *__restrict__ src, int64_t n16)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = __ldg(src + i);
        asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_dst + i), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                     : "memory");
    }
}
}  // namespace ofx

extern "C" int ofx_multicast_copy(void *mc_dst, const void *src, int64_t bytes, int32_t ctas, void *stream)
{
    OFX_REQUIRE(mc_dst && src && bytes >= 0 && bytes % 16 == 0, "ofx_multicast_copy: bad arguments");
    OFX_REQUIRE((((uintptr_t)mc_dst | (uintptr_t)src) & 15) == 0, "ofx_multicast_copy: pointers must be 16-byte aligned");
    if (bytes == 0) return 0;
    if (ctas < 1) ctas = 64;
    ofx::multicast_copy_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>((float4 *)mc_dst, (const float4 *)src, bytes / 16);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_peer_copy(void *dst, const void *src, int64_t bytes, int32_t ctas, void *stream)
{
    OFX_REQUIRE(dst && src && bytes >= 0 && bytes % 16 == 0, "ofx_peer_copy: bad arguments");
    OFX_REQUIRE((((uintptr_t)dst | (uintptr_t)src) & 15) == 0, "ofx_peer_copy: pointers must be 16-byte aligned");
    if (bytes == 0) return 0;
    if (ctas < 1) ctas = 32;
    ofx::peer_copy_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>((uint4 *)dst, (const uint4 *)src, bytes / 16);
    OFX_LAUNCH_CHECK();
    return 0;
}
