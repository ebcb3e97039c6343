// ofx_graph.cu -- graph maintenance and surface-mask operators of libofx.so (sm_100a): the callers either side of
// the fusion hot path when nodes are inserted.
//   ofx_nearest_node      source_frame_kdtree.query(points, k=1)   fusion_with_occlusion/warpfield.py:463
//   ofx_greedy_insert     EDGraph.update, greedy coverage rule      fusion_with_occlusion/embedded_deformation_graph.py:516-536
//   ofx_sample_nodes      sample_nodes (randomShuffle = false)      csrc/cpu/graph_proc.cpp:79-136
//   ofx_erode_mesh        erode_mesh                                csrc/cpu/graph_proc.cpp:17-77
//   ofx_truncated_region  TSDFVolume.compute_truncated_region       fusion_with_occlusion/tsdf.py:706-745
// The greedy rules are order-dependent in the reference (a serial loop over candidates).  Here they run as
//   1. a grid-wide pre-filter (candidates against the existing nodes / the erosion mask) that leaves one bit per
//      candidate, and
//   2. one 1024-thread CTA that repeats "accept the first live candidate, kill every later live candidate it
//      covers" -- the serial loop's result exactly, in as many rounds as nodes are accepted, each round a sweep of
//      the live bits (32 candidates per word, a thread owns its words, so no atomics and no ordering hazards).
// Distances use the float32 operation order of the code they replace (numpy: ((dx^2+dy^2)+dz^2), Eigen's unrolled
// squaredNorm: dx^2+(dy^2+dz^2)); the library is compiled with --fmad=false.
#include <climits>
#include "../../include/ofx.h"
#include "ofx_common.h"
#include "ofx_math.cuh"

namespace ofx {

int sm_count();   // ofx_points.cu

static inline int graph_grid_for(int64_t n, int block, int ctas_per_sm)
{
    int64_t need = (n + block - 1) / block;
    int64_t cap = (int64_t)sm_count() * ctas_per_sm;
    return (int)(need < 1 ? 1 : (need < cap ? need : cap));
}

constexpr int GRAPH_NODE_TILE = 2048;   // 24 KB of node positions per stage
constexpr int GREEDY_THREADS = 1024;

// ---- nearest node ----------------------------------------------------------------------------------------------
// One thread per point, node positions staged tile by tile in shared memory; strict < keeps the lower node index on
// ties.  dist = sqrtf(min d2) (IEEE), as the float32 KD-tree returns it.
__global__ void __launch_bounds__(256) nearest_node_kernel(const float *__restrict__ pts, int64_t n,
                                                           const float *__restrict__ nodes, int M,
                                                           float *__restrict__ dist, int32_t *__restrict__ idx)
{
    __shared__ float s_nodes[GRAPH_NODE_TILE * 3];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t n_round = ((n + stride - 1) / stride) * stride;   // whole CTAs stay in the loop for __syncthreads
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool live = i < n;
        float px = 0.f, py = 0.f, pz = 0.f;
        if (live) { px = __ldg(pts + 3 * i); py = __ldg(pts + 3 * i + 1); pz = __ldg(pts + 3 * i + 2); }
        float best = __int_as_float(0x7f800000);
        int bi = -1;
        for (int base = 0; base < M; base += GRAPH_NODE_TILE) {
            const int cnt = min(GRAPH_NODE_TILE, M - base);
            __syncthreads();
            for (int t = threadIdx.x; t < cnt * 3; t += blockDim.x) s_nodes[t] = __ldg(nodes + (int64_t)base * 3 + t);
            __syncthreads();
            if (live) {
                for (int j = 0; j < cnt; ++j) {
                    const float d2 = dist2(px, py, pz, s_nodes[3 * j], s_nodes[3 * j + 1], s_nodes[3 * j + 2]);
                    if (d2 < best) { best = d2; bi = base + j; }
                }
            }
        }
        if (live) {
            if (dist) dist[i] = sqrtf(best);
            if (idx) idx[i] = bi;
        }
    }
}

// ---- greedy selection: pre-filter --------------------------------------------------------------------------------
// Bit i of `bits` = candidate i may still become a node:
//   EDGraph.update   (mask == NULL): no existing node closer than node_coverage          edg.py:522
//   sample_nodes     (M == 0):       mask[vertex] != 0  (nonErodedVertices)              graph_proc.cpp:115
// Lane l of a warp holds candidate 32*word + l, so one ballot is one word.
__global__ void __launch_bounds__(256) greedy_prefilter_kernel(const float *__restrict__ verts,
                                                               const int32_t *__restrict__ cand, int n,
                                                               const float *__restrict__ nodes, int M, double coverage,
                                                               const uint8_t *__restrict__ mask,
                                                               uint32_t *__restrict__ bits)
{
    __shared__ float s_nodes[GRAPH_NODE_TILE * 3];
    const int stride = gridDim.x * blockDim.x;                       // a multiple of 32
    const int n_round = ((n + stride - 1) / stride) * stride;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool live = i < n;
        const int v = live ? (cand ? __ldg(cand + i) : i) : 0;
        float px = 0.f, py = 0.f, pz = 0.f;
        if (live && M > 0) { px = __ldg(verts + 3 * (int64_t)v); py = __ldg(verts + 3 * (int64_t)v + 1); pz = __ldg(verts + 3 * (int64_t)v + 2); }
        float best = __int_as_float(0x7f800000);
        for (int base = 0; base < M; base += GRAPH_NODE_TILE) {
            const int cnt = min(GRAPH_NODE_TILE, M - base);
            __syncthreads();
            for (int t = threadIdx.x; t < cnt * 3; t += blockDim.x) s_nodes[t] = __ldg(nodes + (int64_t)base * 3 + t);
            __syncthreads();
            if (live)
                for (int j = 0; j < cnt; ++j)
                    best = fminf(best, dist2(s_nodes[3 * j], s_nodes[3 * j + 1], s_nodes[3 * j + 2], px, py, pz));   // self.nodes - v
        }
        bool alive = live;
        if (live && M > 0) alive = !((double)sqrtf(best) < coverage);      // np.min(np.linalg.norm(...)) < node_coverage
        if (live && mask) alive = alive && __ldg(mask + v) != 0;
        const uint32_t word = __ballot_sync(0xffffffffu, alive);
        if ((threadIdx.x & 31) == 0 && i < n) bits[i >> 5] = word;
    }
}

// ---- greedy selection: the serial rule, one CTA ------------------------------------------------------------------
__device__ __forceinline__ int block_min(int v, int *s_red)
{
    v = __reduce_min_sync(0xffffffffu, v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    int r = lane < nwarps ? s_red[lane] : INT_MAX;
    r = __reduce_min_sync(0xffffffffu, r);
    __syncthreads();     // s_red is reused by the next call
    return r;
}

// EIGEN_ORDER = false: (double)sqrtf(((dx^2+dy^2)+dz^2)) < coverage            edg.py:530-531 (numpy float32 norm)
// EIGEN_ORDER = true : dx^2+(dy^2+dz^2) <= coverage^2 (float32)                graph_proc.cpp:89,121 (Eigen squaredNorm)
template <bool EIGEN_ORDER>
__global__ void __launch_bounds__(GREEDY_THREADS) greedy_select_kernel(const float *__restrict__ verts,
                                                                       const int32_t *__restrict__ cand, int n,
                                                                       uint32_t *__restrict__ bits, int nwords,
                                                                       double coverage, float coverage2,
                                                                       int32_t *__restrict__ accepted,
                                                                       float *__restrict__ accepted_pos,
                                                                       int32_t *__restrict__ n_accepted)
{
    __shared__ int s_red[32];
    const int tid = threadIdx.x;
    int my_min = INT_MAX;
    for (int w = tid; w < nwords; w += GREEDY_THREADS) {
        const uint32_t m = bits[w];
        if (m) { my_min = w * 32 + __ffs(m) - 1; break; }
    }
    int head = block_min(my_min, s_red);
    int n_acc = 0;
    while (head < n) {                                               // head is uniform across the CTA
        const int hv = cand ? __ldg(cand + head) : head;
        const float ax = __ldg(verts + 3 * (int64_t)hv), ay = __ldg(verts + 3 * (int64_t)hv + 1), az = __ldg(verts + 3 * (int64_t)hv + 2);
        if (tid == 0) {
            accepted[n_acc] = hv;
            if (accepted_pos) { accepted_pos[3 * n_acc] = ax; accepted_pos[3 * n_acc + 1] = ay; accepted_pos[3 * n_acc + 2] = az; }
        }
        ++n_acc;
        const int hw = head >> 5;
        int w = tid;
        if (w < hw) w += ((hw - w + GREEDY_THREADS - 1) / GREEDY_THREADS) * GREEDY_THREADS;   // words before head's hold no live bit
        my_min = INT_MAX;
        for (; w < nwords; w += GREEDY_THREADS) {
            uint32_t m = bits[w];
            if (w == hw) m &= ~((2u << (head & 31)) - 1u);           // drop the accepted candidate (and the dead bits below it)
            uint32_t keep = m;
            while (m) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                const int i = w * 32 + b;
                const int v = cand ? __ldg(cand + i) : i;
                const float dx = __ldg(verts + 3 * (int64_t)v) - ax, dy = __ldg(verts + 3 * (int64_t)v + 1) - ay,
                            dz = __ldg(verts + 3 * (int64_t)v + 2) - az;
                bool covered;
                if (EIGEN_ORDER) covered = ((dx * dx) + ((dy * dy) + (dz * dz))) <= coverage2;
                else covered = (double)sqrtf(((dx * dx) + (dy * dy)) + (dz * dz)) < coverage;
                if (covered) keep &= ~(1u << b);
            }
            bits[w] = keep;
            if (keep && my_min == INT_MAX) my_min = w * 32 + __ffs(keep) - 1;
        }
        head = block_min(my_min, s_red);
    }
    if (tid == 0) *n_accepted = n_acc;
}

// ---- erode_mesh ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool face_ok(const int32_t *__restrict__ faces, int f, int V, int &a, int &b, int &c)
{
    a = __ldg(faces + 3 * (int64_t)f); b = __ldg(faces + 3 * (int64_t)f + 1); c = __ldg(faces + 3 * (int64_t)f + 2);
    return (unsigned)a < (unsigned)V && (unsigned)b < (unsigned)V && (unsigned)c < (unsigned)V;   // out-of-range faces are ignored
}

__global__ void __launch_bounds__(256) erode_count_kernel(const int32_t *__restrict__ faces, const uint8_t *__restrict__ alive,
                                                          int F, int V, int *__restrict__ cnt)
{
    for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < F; f += gridDim.x * blockDim.x) {
        int a, b, c;
        if (!alive[f] || !face_ok(faces, f, V, a, b, c)) continue;
        atomicAdd(cnt + a, 1); atomicAdd(cnt + b, 1); atomicAdd(cnt + c, 1);                       // graph_proc.cpp:39-44
    }
}

__global__ void __launch_bounds__(256) erode_filter_kernel(const int32_t *__restrict__ faces, uint8_t *__restrict__ alive, int F,
                                                           int V, const int *__restrict__ cnt, int min_neighbours)
{
    for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < F; f += gridDim.x * blockDim.x) {
        int a, b, c;
        if (!alive[f]) continue;
        if (!face_ok(faces, f, V, a, b, c)) { alive[f] = 0; continue; }
        if (!(cnt[a] >= min_neighbours && cnt[b] >= min_neighbours && cnt[c] >= min_neighbours)) alive[f] = 0;   // :49-54
    }
}

__global__ void __launch_bounds__(256) erode_mark_kernel(const int32_t *__restrict__ faces, const uint8_t *__restrict__ alive, int F,
                                                         int V, uint8_t *__restrict__ non_eroded)
{
    for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < F; f += gridDim.x * blockDim.x) {
        int a, b, c;
        if (!alive[f] || !face_ok(faces, f, V, a, b, c)) continue;
        non_eroded[a] = 1; non_eroded[b] = 1; non_eroded[c] = 1;                                    // :65-70
    }
}

// ---- compute_truncated_region -------------------------------------------------------------------------------------
// Streaming: 4 B read + 1 B written per voxel; only voxels with |tsdf| <= 0.9 (the band) look at their 26 neighbours
// (L1/L2 hits: neighbouring rows).  VEC = 4: one float4 load and one 32-bit store per thread (Dz % 4 == 0).
__device__ __forceinline__ uint8_t truncated_voxel(const float *__restrict__ t, float c, int w, int h, int d, int W, int H, int D,
                                                   double max_diff)
{
    if ((double)fabsf(c) > 0.9) return 0;                                                           // tsdf.py:724
    if (d == 0 || d == D - 1 || h == 0 || h == H - 1 || w == 0 || w == W - 1) return 0;            // :729
    for (int dw = -1; dw < 2; ++dw)
        for (int dh = -1; dh < 2; ++dh) {
            const float *row = t + ((int64_t)(w + dw) * H + (h + dh)) * D + d;
#pragma unroll
            for (int dd = -1; dd < 2; ++dd)
                if ((double)fabsf(__ldg(row + dd) - c) > max_diff) return 0;                        // :740
        }
    return 1;
}

template <int VEC>
__global__ void __launch_bounds__(256) truncated_region_kernel(const float *__restrict__ t, int W, int H, int D, double max_diff,
                                                               uint8_t *__restrict__ out)
{
    const int64_t n = (int64_t)W * H * D / VEC;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t base = i * VEC;
        const int d0 = (int)(base % D);
        const int h = (int)((base / D) % H);
        const int w = (int)(base / ((int64_t)D * H));
        if (VEC == 4) {
            const float4 c = __ldg(reinterpret_cast<const float4 *>(t + base));
            uchar4 r;
            r.x = truncated_voxel(t, c.x, w, h, d0, W, H, D, max_diff);
            r.y = truncated_voxel(t, c.y, w, h, d0 + 1, W, H, D, max_diff);
            r.z = truncated_voxel(t, c.z, w, h, d0 + 2, W, H, D, max_diff);
            r.w = truncated_voxel(t, c.w, w, h, d0 + 3, W, H, D, max_diff);
            *reinterpret_cast<uchar4 *>(out + base) = r;
        } else {
            out[base] = truncated_voxel(t, __ldg(t + base), w, h, d0, W, H, D, max_diff);
        }
    }
}

static int run_greedy(bool eigen_order, const float *verts, const int32_t *cand, int n, const float *nodes, int M,
                      double coverage, float coverage2, const uint8_t *mask, int32_t *accepted, float *accepted_pos,
                      int32_t *n_accepted, cudaStream_t s)
{
    if (n == 0) {
        OFX_CUDA(cudaMemsetAsync(n_accepted, 0, sizeof(int32_t), s));
        return 0;
    }
    const int nwords = (n + 31) / 32;
    uint32_t *bits = nullptr;
    OFX_CUDA(cudaMallocAsync((void **)&bits, (size_t)nwords * 4, s));
    greedy_prefilter_kernel<<<graph_grid_for(n, 256, 4), 256, 0, s>>>(verts, cand, n, nodes, M, coverage, mask, bits);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) {
        if (eigen_order)
            greedy_select_kernel<true><<<1, GREEDY_THREADS, 0, s>>>(verts, cand, n, bits, nwords, coverage, coverage2, accepted,
                                                                   accepted_pos, n_accepted);
        else
            greedy_select_kernel<false><<<1, GREEDY_THREADS, 0, s>>>(verts, cand, n, bits, nwords, coverage, coverage2, accepted,
                                                                    accepted_pos, n_accepted);
        e = cudaGetLastError();
    }
    cudaFreeAsync(bits, s);
    if (e != cudaSuccess) return fail("%s:%d: greedy kernels -> %s", __FILE__, __LINE__, cudaGetErrorString(e));
    return 0;
}

}  // namespace ofx

using namespace ofx;

extern "C" int ofx_nearest_node(const float *points, int64_t n, const float *nodes, int32_t M, float *dist, int32_t *idx,
                                void *stream)
{
    OFX_REQUIRE(n >= 0 && M >= 1, "ofx_nearest_node: need n >= 0 and M >= 1 (n=%lld, M=%d)", (long long)n, M);
    if (n == 0) return 0;
    OFX_REQUIRE(points && nodes && (dist || idx), "ofx_nearest_node: null pointer");
    nearest_node_kernel<<<graph_grid_for(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(points, n, nodes, M, dist, idx);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_greedy_insert(const float *verts, int32_t V, const int32_t *cand, int32_t n_cand, const float *nodes,
                                 int32_t M, double node_coverage, int32_t *accepted, int32_t *n_accepted, void *stream)
{
    OFX_REQUIRE(V >= 0 && n_cand >= 0 && M >= 0 && n_accepted, "ofx_greedy_insert: bad arguments");
    OFX_REQUIRE(n_cand == 0 || (verts && cand && accepted && (M == 0 || nodes)), "ofx_greedy_insert: null pointer");
    return run_greedy(false, verts, cand, n_cand, nodes, M, node_coverage, 0.0f, nullptr, accepted, nullptr, n_accepted,
                      (cudaStream_t)stream);
}

extern "C" int ofx_sample_nodes(const float *verts, int32_t V, const uint8_t *non_eroded, float node_coverage,
                                int32_t use_only_non_eroded, int32_t *node_idx, float *node_pos, int32_t *n_nodes,
                                void *stream)
{
    OFX_REQUIRE(V >= 0 && n_nodes, "ofx_sample_nodes: bad arguments");
    OFX_REQUIRE(V == 0 || (verts && node_idx && (!use_only_non_eroded || non_eroded)), "ofx_sample_nodes: null pointer");
    const float c2 = node_coverage * node_coverage;                                                 // graph_proc.cpp:89
    return run_greedy(true, verts, nullptr, V, nullptr, 0, 0.0, c2, use_only_non_eroded ? non_eroded : nullptr, node_idx,
                      node_pos, n_nodes, (cudaStream_t)stream);
}

extern "C" int ofx_erode_mesh(const int32_t *faces, int32_t F, int32_t V, int32_t n_iterations, int32_t min_neighbours,
                              uint8_t *non_eroded, void *stream)
{
    OFX_REQUIRE(F >= 0 && V >= 0, "ofx_erode_mesh: bad arguments");
    if (V == 0) return 0;
    OFX_REQUIRE(non_eroded && (F == 0 || faces), "ofx_erode_mesh: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    OFX_CUDA(cudaMemsetAsync(non_eroded, 0, (size_t)V, s));
    if (F == 0) return 0;
    uint8_t *alive = nullptr;
    int *cnt = nullptr;
    OFX_CUDA(cudaMallocAsync((void **)&alive, (size_t)F, s));
    cudaError_t e = cudaMallocAsync((void **)&cnt, (size_t)V * sizeof(int), s);
    if (e != cudaSuccess) { cudaFreeAsync(alive, s); return fail("ofx_erode_mesh: %s", cudaGetErrorString(e)); }
    const int grid = graph_grid_for(F, 256, 8);
    e = cudaMemsetAsync(alive, 1, (size_t)F, s);
    for (int it = 0; it < n_iterations && e == cudaSuccess; ++it) {
        e = cudaMemsetAsync(cnt, 0, (size_t)V * sizeof(int), s);
        if (e != cudaSuccess) break;
        erode_count_kernel<<<grid, 256, 0, s>>>(faces, alive, F, V, cnt);
        erode_filter_kernel<<<grid, 256, 0, s>>>(faces, alive, F, V, cnt, min_neighbours);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) {
        erode_mark_kernel<<<grid, 256, 0, s>>>(faces, alive, F, V, non_eroded);
        e = cudaGetLastError();
    }
    cudaFreeAsync(alive, s);
    cudaFreeAsync(cnt, s);
    if (e != cudaSuccess) return fail("%s:%d: erode kernels -> %s", __FILE__, __LINE__, cudaGetErrorString(e));
    return 0;
}

extern "C" int ofx_truncated_region(const float *tsdf, int32_t Dx, int32_t Dy, int32_t Dz, double max_diff, uint8_t *region,
                                    void *stream)
{
    OFX_REQUIRE(Dx >= 0 && Dy >= 0 && Dz >= 0, "ofx_truncated_region: bad arguments");
    const int64_t n = (int64_t)Dx * Dy * Dz;
    if (n == 0) return 0;
    OFX_REQUIRE(tsdf && region, "ofx_truncated_region: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    const bool vec = Dz % 4 == 0 && (((uintptr_t)tsdf & 15) == 0) && (((uintptr_t)region & 3) == 0);
    if (vec) truncated_region_kernel<4><<<graph_grid_for(n / 4, 256, 8), 256, 0, s>>>(tsdf, Dx, Dy, Dz, max_diff, region);
    else truncated_region_kernel<1><<<graph_grid_for(n, 256, 8), 256, 0, s>>>(tsdf, Dx, Dy, Dz, max_diff, region);
    OFX_LAUNCH_CHECK();
    return 0;
}
