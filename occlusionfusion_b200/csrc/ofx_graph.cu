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
//      covers" -- the serial loop's result exactly, in as many rounds as nodes are accepted; a round sweeps only the
//      candidates binned into the 27 hash cells around the accepted vertex (v1 swept every live bit: 59 / 322 ms for
//      90 k / 490 k vertices).
// Distances use the float32 operation order of the code they replace (numpy: ((dx^2+dy^2)+dz^2), Eigen's unrolled
// squaredNorm: dx^2+(dy^2+dz^2)); the library is compiled with --fmad=false.
#include <climits>
#include <cmath>
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

// Scratch for the graph operators comes from a stream-ordered pool of our own (one per device) that keeps its memory
// between calls: the default pool hands memory back at every synchronisation, which made each call pay milliseconds of
// allocation (first measurement: erode_mesh 11 ms, 3 ms of it kernels).
static cudaMemPool_t g_pool[64] = {};
static int scratch_alloc(void **p, size_t bytes, cudaStream_t s)
{
    int dev = 0;
    OFX_CUDA(cudaGetDevice(&dev));
    OFX_REQUIRE(dev >= 0 && dev < 64, "scratch_alloc: device %d out of range", dev);
    if (!g_pool[dev]) {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        OFX_CUDA(cudaMemPoolCreate(&g_pool[dev], &props));
        uint64_t keep = UINT64_MAX;
        OFX_CUDA(cudaMemPoolSetAttribute(g_pool[dev], cudaMemPoolAttrReleaseThreshold, &keep));
    }
    OFX_CUDA(cudaMallocFromPoolAsync(p, bytes < 16 ? 16 : bytes, g_pool[dev], s));
    return 0;
}

int graph_scratch_alloc(void **p, size_t bytes, cudaStream_t s) { return scratch_alloc(p, bytes, s); }   // for ofx_graph_edges.cu

// largest float a with (double)a <= x: for every float v, ((double)v > x) == (v > a)
static float float_at_most(double x)
{
    float a = (float)x;
    if ((double)a > x) a = nextafterf(a, -INFINITY);
    return a;
}

constexpr int GRAPH_NODE_TILE = 2048;   // 24 KB of node positions per stage
constexpr int GREEDY_THREADS = 1024;

// ---- nearest node ----------------------------------------------------------------------------------------------
// One thread per point, node positions staged tile by tile in shared memory; strict < keeps the lower node index on
// ties.  dist = sqrtf(min d2) (IEEE), as the float32 KD-tree returns it.
// BY_NORM: compare the float32 norms instead of the squared distances -- np.argmin(np.linalg.norm(pts - q, axis=1)) keeps the
// FIRST index among equal norms, and two different squared distances can round to the same norm (edg.py:545-547).
template <bool BY_NORM>
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
                    float d2 = dist2(px, py, pz, s_nodes[3 * j], s_nodes[3 * j + 1], s_nodes[3 * j + 2]);
                    if (BY_NORM) d2 = sqrtf(d2);
                    if (d2 < best) { best = d2; bi = base + j; }
                }
            }
        }
        if (live) {
            if (dist) dist[i] = BY_NORM ? best : sqrtf(best);
            if (idx) idx[i] = bi;
        }
    }
}

// np.argmin(np.linalg.norm(points - q, axis=1)) for FEW queries against MANY points (EDGraph.update re-anchors a few hundred
// nodes on a mesh of 10^4..10^6 vertices): one CTA per query, the threads stride over the points, (norm, index) pairs are
// reduced keeping the first index among equal norms.  (One thread per query took 1.1 ms for 381 x 14.5 k.)
__global__ void __launch_bounds__(256) argmin_norm_wide_kernel(const float *__restrict__ queries, const float *__restrict__ pts, int M,
                                                               float *__restrict__ dist, int32_t *__restrict__ idx)
{
    __shared__ float s_d[8];
    __shared__ int s_i[8];
    const int q = blockIdx.x;
    const float qx = __ldg(queries + 3 * q), qy = __ldg(queries + 3 * q + 1), qz = __ldg(queries + 3 * q + 2);
    float best = __int_as_float(0x7f800000);
    int bi = 0x7fffffff;
    for (int j = threadIdx.x; j < M; j += blockDim.x) {
        const float d = sqrtf(dist2(qx, qy, qz, __ldg(pts + 3 * (int64_t)j), __ldg(pts + 3 * (int64_t)j + 1), __ldg(pts + 3 * (int64_t)j + 2)));
        if (d < best) { best = d; bi = j; }              // ascending j per thread: the first index among equal norms
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float od = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (od < best || (od == best && oi < bi)) { best = od; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { s_d[threadIdx.x >> 5] = best; s_i[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w)
            if (s_d[w] < best || (s_d[w] == best && s_i[w] < bi)) { best = s_d[w]; bi = s_i[w]; }
        // (dist2 is the squared distance with numpy's summation order; NaN points never win, as in np.argmin... of a row
        // without NaN: meshes carry none)
        if (dist) dist[q] = best;
        if (idx) idx[q] = bi == 0x7fffffff ? 0 : bi;
    }
}

// ---- pixel anchors (Euclidean) -------------------------------------------------------------------------------------
// compute_pixel_anchors_euclidean (csrc/cpu/graph_proc.cpp:610-709): for every pixel of a (3,H,W) point image with z > 0 the
// PIXEL_K = 4 nearest nodes by squared distance (Eigen order dx^2 + (dy^2 + dz^2)), kept sorted by insertion with <=, so that
// among equal distances the later node comes first (:656), Gaussian weights exp(-d2 / (2 c^2)) normalised by their sum (no
// 1e-6 here, :694-699).  One thread per pixel, node tiles in shared memory.  exp is evaluated in float64 and rounded once,
// which is what glibc's float exp amounts to in all but ~1e-8 of the cases (the reference calls std::exp on a float).
constexpr int PIXEL_K = 4;
__global__ void __launch_bounds__(256) pixel_anchors_kernel(const float *__restrict__ nodes, int M, const float *__restrict__ img,
                                                            int64_t npix, float coverage, int32_t *__restrict__ anchors,
                                                            float *__restrict__ weights)
{
    __shared__ float s_nodes[GRAPH_NODE_TILE * 3];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t n_round = ((npix + stride - 1) / stride) * stride;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool live = i < npix;
        float px = 0.f, py = 0.f, pz = 0.f;
        if (live) { px = __ldg(img + i); py = __ldg(img + npix + i); pz = __ldg(img + 2 * npix + i); }
        const bool use = live && pz > 0.0f;                                                   // :641
        float bd[PIXEL_K];
        int bi[PIXEL_K];
        int cnt = 0;
#pragma unroll
        for (int k = 0; k < PIXEL_K; ++k) { bd[k] = 0.f; bi[k] = -1; }
        for (int base = 0; base < M; base += GRAPH_NODE_TILE) {
            const int tile = min(GRAPH_NODE_TILE, M - base);
            __syncthreads();
            for (int t = threadIdx.x; t < tile * 3; t += blockDim.x) s_nodes[t] = __ldg(nodes + (int64_t)base * 3 + t);
            __syncthreads();
            if (!use) continue;
            for (int j = 0; j < tile; ++j) {
                const float dx = px - s_nodes[3 * j], dy = py - s_nodes[3 * j + 1], dz = pz - s_nodes[3 * j + 2];
                const float d2 = (dx * dx) + ((dy * dy) + (dz * dz));
                // first position whose distance is >= d2 (:656); none: append while fewer than K are held (:663)
                int pos = cnt;
#pragma unroll
                for (int k = PIXEL_K - 1; k >= 0; --k)
                    if (k < cnt && d2 <= bd[k]) pos = k;
                if (pos >= PIXEL_K) continue;
#pragma unroll
                for (int k = PIXEL_K - 1; k > 0; --k)
                    if (k > pos) { bd[k] = bd[k - 1]; bi[k] = bi[k - 1]; }
#pragma unroll
                for (int k = 0; k < PIXEL_K; ++k)
                    if (k == pos) { bd[k] = d2; bi[k] = base + j; }
                if (cnt < PIXEL_K) ++cnt;
            }
        }
        if (!live) continue;
        float w[PIXEL_K], sum = 0.0f;
        const float denom = 2.f * coverage * coverage;                                        // :148
#pragma unroll
        for (int k = 0; k < PIXEL_K; ++k) {
            w[k] = k < cnt ? (float)exp((double)(-bd[k] / denom)) : 0.0f;
            if (k < cnt) sum += w[k];
        }
#pragma unroll
        for (int k = 0; k < PIXEL_K; ++k) {
            float v = 0.0f;
            if (k < cnt) v = sum > 0.0f ? w[k] / sum : 1.f / (float)cnt;                      // :694-699
            anchors[PIXEL_K * i + k] = k < cnt ? bi[k] : -1;
            weights[PIXEL_K * i + k] = v;
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

// ---- greedy selection: the serial rule, one CTA over a spatial hash ------------------------------------------------
// An accepted vertex can only cover candidates closer than the coverage radius, so the live candidates are binned into
// cells a little wider than that radius (hashed into GREEDY_CELLS slots, counting sort; positions copied next to the ids
// in cell order so that a sweep reads contiguous memory) and a round looks at the 27 cells around the accepted vertex
// instead of at every live candidate.  The next head is found by scanning the live words forward from the previous head
// (the head only moves forward: one pass over the words in total).  Slot collisions only add candidates to a sweep; the
// distance test decides.
constexpr int GREEDY_CELLS = 1 << 14;   // slots; collisions only lengthen a sweep (v2 used 2^16: the scan alone took 110 us)

__device__ __forceinline__ int cell_coord(float p, double inv_cell)
{
    double x = floor((double)p * inv_cell);
    if (!(x == x)) x = 0.0;
    x = fmin(fmax(x, -1073741824.0), 1073741824.0);
    return (int)x;
}

__device__ __forceinline__ uint32_t cell_slot(int ix, int iy, int iz)
{
    return ((uint32_t)ix * 73856093u ^ (uint32_t)iy * 19349663u ^ (uint32_t)iz * 83492791u) & (GREEDY_CELLS - 1);
}

__global__ void __launch_bounds__(256) cell_count_kernel(const float *__restrict__ verts, const int32_t *__restrict__ cand, int n,
                                                         const uint32_t *__restrict__ bits, double inv_cell,
                                                         uint32_t *__restrict__ slot_of, int *__restrict__ count)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if (!((bits[i >> 5] >> (i & 31)) & 1u)) continue;
        const int v = cand ? __ldg(cand + i) : i;
        const uint32_t h = cell_slot(cell_coord(__ldg(verts + 3 * (int64_t)v), inv_cell), cell_coord(__ldg(verts + 3 * (int64_t)v + 1), inv_cell),
                                     cell_coord(__ldg(verts + 3 * (int64_t)v + 2), inv_cell));
        slot_of[i] = h;
        atomicAdd(count + h, 1);
    }
}

// exclusive scan of count[GREEDY_CELLS] -> start[GREEDY_CELLS + 1], cursor = start; one CTA of 1024 threads
__global__ void __launch_bounds__(1024) cell_scan_kernel(const int *__restrict__ count, int *__restrict__ start, int *__restrict__ cursor)
{
    constexpr int PER = GREEDY_CELLS / 1024;
    __shared__ int s_warp[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int local[PER];
    int sum = 0;
#pragma unroll
    for (int k = 0; k < PER; k += 4) {
        const int4 c4 = *reinterpret_cast<const int4 *>(count + tid * PER + k);
        local[k] = c4.x; local[k + 1] = c4.y; local[k + 2] = c4.z; local[k + 3] = c4.w;
        sum += (c4.x + c4.y) + (c4.z + c4.w);
    }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
        s_warp[lane] = w;
    }
    __syncthreads();
    int run = incl - sum + (warp ? s_warp[warp - 1] : 0);
#pragma unroll
    for (int k = 0; k < PER; ++k) { start[tid * PER + k] = run; cursor[tid * PER + k] = run; run += local[k]; }
    if (tid == 1023) start[GREEDY_CELLS] = run;
}

__global__ void __launch_bounds__(256) cell_scatter_kernel(const float *__restrict__ verts, const int32_t *__restrict__ cand, int n,
                                                           const uint32_t *__restrict__ bits, const uint32_t *__restrict__ slot_of,
                                                           int *__restrict__ cursor, int *__restrict__ sidx, float *__restrict__ sx,
                                                           float *__restrict__ sy, float *__restrict__ sz)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if (!((bits[i >> 5] >> (i & 31)) & 1u)) continue;
        const int v = cand ? __ldg(cand + i) : i;
        const int pos = atomicAdd(cursor + slot_of[i], 1);
        sidx[pos] = i;
        sx[pos] = __ldg(verts + 3 * (int64_t)v); sy[pos] = __ldg(verts + 3 * (int64_t)v + 1); sz[pos] = __ldg(verts + 3 * (int64_t)v + 2);
    }
}

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
                                                                       uint32_t *bits, int nwords, double coverage,
                                                                       float coverage2, double inv_cell,
                                                                       const int *__restrict__ start, const int *__restrict__ sidx,
                                                                       const float *__restrict__ sx, const float *__restrict__ sy,
                                                                       const float *__restrict__ sz,
                                                                       int32_t *__restrict__ accepted,
                                                                       float *__restrict__ accepted_pos,
                                                                       int32_t *__restrict__ n_accepted)
{
    __shared__ int s_red[32];
    __shared__ int s_begin[27], s_prefix[28];
    const int tid = threadIdx.x;
    int n_acc = 0;
    int wpos = 0;                                                    // words before wpos hold no live bit (uniform)
    for (;;) {
        // ---- next head: first live bit at or after word wpos (bits are read around L1: other threads clear them with atomics)
        int head = INT_MAX;
        while (wpos < nwords) {
            const int w = wpos + tid;
            const uint32_t m = w < nwords ? __ldcg(bits + w) : 0u;
            head = block_min(m ? w * 32 + __ffs(m) - 1 : INT_MAX, s_red);
            if (head != INT_MAX) { wpos = head >> 5; break; }
            wpos += GREEDY_THREADS;
        }
        if (head == INT_MAX) break;
        // ---- accept it
        const int hv = cand ? __ldg(cand + head) : head;
        const float ax = __ldg(verts + 3 * (int64_t)hv), ay = __ldg(verts + 3 * (int64_t)hv + 1), az = __ldg(verts + 3 * (int64_t)hv + 2);
        if (tid == 0) {
            accepted[n_acc] = hv;
            if (accepted_pos) { accepted_pos[3 * n_acc] = ax; accepted_pos[3 * n_acc + 1] = ay; accepted_pos[3 * n_acc + 2] = az; }
            atomicAnd(bits + (head >> 5), ~(1u << (head & 31)));
        }
        ++n_acc;
        // ---- the 27 cells around it
        if (tid < 27) {
            const int cx = cell_coord(ax, inv_cell) + tid % 3 - 1, cy = cell_coord(ay, inv_cell) + (tid / 3) % 3 - 1,
                      cz = cell_coord(az, inv_cell) + tid / 9 - 1;
            const uint32_t h = cell_slot(cx, cy, cz);
            const int b0 = __ldg(start + h), len = __ldg(start + h + 1) - b0;
            s_begin[tid] = b0;
            // inclusive prefix of the 27 lengths inside warp 0
            int incl = len;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0x07ffffffu, incl, o); if (tid >= o) incl += y; }
            s_prefix[tid + 1] = incl;
            if (tid == 0) s_prefix[0] = 0;
        }
        __syncthreads();
        const int total = s_prefix[27];
        for (int f = tid; f < total; f += GREEDY_THREADS) {
            int k = 0;
#pragma unroll
            for (int q = 1; q < 27; ++q) k += (f >= s_prefix[q]);
            const int j = s_begin[k] + (f - s_prefix[k]);
            const int i = __ldg(sidx + j);
            const float dx = __ldg(sx + j) - ax, dy = __ldg(sy + j) - ay, dz = __ldg(sz + j) - az;
            bool covered;
            if (EIGEN_ORDER) covered = ((dx * dx) + ((dy * dy) + (dz * dz))) <= coverage2;
            else covered = (double)sqrtf(((dx * dx) + (dy * dy)) + (dz * dz)) < coverage;
            if (covered && i > head) atomicAnd(bits + (i >> 5), ~(1u << (i & 31)));
        }
        __syncthreads();                                             // the kills are visible to the next head search
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
// Streaming: 4 B read + 1 B written per voxel.  The float64 comparisons of the reference ((double)|x| > 0.9,
// (double)|nb - c| > max_diff) are done in float32 against the largest float <= the bound (same truth value for every
// float).  "Some neighbour differs by more than max_diff" is decided from the min and max over the 3x3x3 block:
// fl(nb - c) is monotone in nb, so the largest |nb - c| is fl(max - c) or fl(c - min).
struct TruncBounds { float band, diff; };
#ifndef OFX_TR_CTAS
#define OFX_TR_CTAS 4          // CTAs per SM in the grid of the vector kernel = the resident count (56-64 registers); 8 and 16 measured 4-8 % slower
#endif
#ifndef OFX_TR_STREAM
#define OFX_TR_STREAM 0        // 1: evict-first loads / stores for the streamed volume (measured: no difference)
#endif   // float_at_most(0.9), float_at_most(max_diff)

// scalar form (any Dz): one thread per voxel
__global__ void __launch_bounds__(256) truncated_region_scalar_kernel(const float *__restrict__ t, int W, int H, int D, TruncBounds b,
                                                                      uint8_t *__restrict__ out)
{
    const int64_t n = (int64_t)W * H * D;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int d = (int)(i % D), h = (int)((i / D) % H), w = (int)(i / ((int64_t)D * H));
        const float c = __ldg(t + i);
        uint8_t r = 0;
        if (!(fabsf(c) > b.band) && d > 0 && d < D - 1 && h > 0 && h < H - 1 && w > 0 && w < W - 1) {   // tsdf.py:724, 729
            float mn = c, mx = c;
            for (int dw = -1; dw < 2; ++dw)
                for (int dh = -1; dh < 2; ++dh) {
                    const float *row = t + i + ((int64_t)dw * H + dh) * D;
#pragma unroll
                    for (int dd = -1; dd < 2; ++dd) { const float v = __ldg(row + dd); mn = fminf(mn, v); mx = fmaxf(mx, v); }
                }
            r = !((mx - c > b.diff) || (c - mn > b.diff));                                               // :740
        }
        out[i] = r;
    }
}

// vector form (Dz % 4 == 0): a warp walks z-rows (row r = w*H + h, consecutive warps of a CTA take consecutive rows, so the
// rows h-1 / h+1 a warp needs are the ones its neighbours just streamed), a lane owns 4 consecutive z voxels of the row at a
// time (one float4, one 32-bit store; the next quad is loaded before the current one is examined).  No 64-bit divisions (v1
// spent its time in them).  Only when some lane of the warp holds a band voxel does the warp look at neighbours: per
// neighbouring row one coalesced float4 load per lane, the z-1 / z+4 values from the adjacent lanes by shuffle (the two edge
// lanes load them); min and max are kept per z-column over the 9 rows and folded over z once at the end.
__global__ void __launch_bounds__(256) truncated_region_vec_kernel(const float *__restrict__ t, int W, int H, int D, TruncBounds b,
                                                                   uint8_t *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const int warps = gridDim.x * (blockDim.x >> 5);
    const int qpr = D >> 2;                                                // quads per row
    const int ipr = (qpr + 31) >> 5;                                       // 32-quad chunks per row
    const int rows = W * H;
    const float INF = __int_as_float(0x7f800000);
    // Work items of this warp: (row, chunk) pairs, rows r0, r0 + warps, ...  Three items of look-ahead, carried ACROSS row
    // boundaries (v2 restarted the look-ahead at every row: half of its warp-state samples sat on that first load).
    int fr = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), fc = 0;  // position of the next fetch
    auto fetch = [&]() -> float4 {
        float4 v = make_float4(1.f, 1.f, 1.f, 1.f);
        const int q = (fc << 5) + lane;
#if OFX_TR_STREAM
        if (fr < rows && q < qpr) v = __ldcs(reinterpret_cast<const float4 *>(t + (int64_t)fr * D) + q);
#else
        if (fr < rows && q < qpr) v = __ldg(reinterpret_cast<const float4 *>(t + (int64_t)fr * D) + q);
#endif
        if (++fc == ipr) { fc = 0; fr += warps; }
        return v;
    };
    int r = fr, ch = 0;                                                    // position of the item being examined
    float4 n1 = fetch(), n2 = fetch(), n3 = fetch();
    while (r < rows) {
        const int w = r / H, h = r - w * H;
        const bool row_in = h > 0 && h < H - 1 && w > 0 && w < W - 1;
        const float *rowp = t + (int64_t)r * D;
        uint8_t *outp = out + (int64_t)r * D;
        for (; ch < ipr; ++ch) {
            const int q = (ch << 5) + lane, d0 = q << 2;
            const bool live = q < qpr;
            const float4 c4 = n1;
            n1 = n2; n2 = n3; n3 = fetch();
            const float c[4] = {c4.x, c4.y, c4.z, c4.w};
            bool cand[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) cand[e] = live && row_in && !(fabsf(c[e]) > b.band) && d0 + e > 0 && d0 + e < D - 1;   // tsdf.py:724, 729
            const bool any = cand[0] || cand[1] || cand[2] || cand[3];
            uchar4 res = make_uchar4(0, 0, 0, 0);
            if (__any_sync(0xffffffffu, any)) {                           // row_in holds for the whole warp here (one row per warp)
                const bool has_l = live && d0 > 0, has_r = live && d0 + 4 < D;
                float mn[6], mx[6];
#pragma unroll
                for (int e = 0; e < 6; ++e) { mn[e] = INF; mx[e] = -INF; }
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    const int dw = k / 3 - 1, dh = k % 3 - 1;
                    const float *nrow = rowp + ((int64_t)dw * H + dh) * D + d0;
                    float4 qv = c4;
                    if (k != 4 && live) qv = __ldg(reinterpret_cast<const float4 *>(nrow));
                    float l = __shfl_up_sync(0xffffffffu, qv.w, 1), rr = __shfl_down_sync(0xffffffffu, qv.x, 1);
                    if (lane == 0 && has_l) l = __ldg(nrow - 1);
                    if (lane == 31 && has_r) rr = __ldg(nrow + 4);
                    if (!has_l) l = qv.x;          // z = 0 / Dz-1 are never candidates: any in-range value will do
                    if (!has_r) rr = qv.w;
                    const float v[6] = {l, qv.x, qv.y, qv.z, qv.w, rr};
#pragma unroll
                    for (int e = 0; e < 6; ++e) { mn[e] = fminf(mn[e], v[e]); mx[e] = fmaxf(mx[e], v[e]); }
                }
                uint8_t rb[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float lo = fminf(mn[e], fminf(mn[e + 1], mn[e + 2])), hi = fmaxf(mx[e], fmaxf(mx[e + 1], mx[e + 2]));
                    rb[e] = cand[e] && !((hi - c[e] > b.diff) || (c[e] - lo > b.diff));                                        // :740
                }
                res = make_uchar4(rb[0], rb[1], rb[2], rb[3]);
            }
#if OFX_TR_STREAM
            if (live) __stcs(reinterpret_cast<uchar4 *>(outp + d0), res);
#else
            if (live) *reinterpret_cast<uchar4 *>(outp + d0) = res;
#endif
        }
        ch = 0;
        r += warps;
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
    // one scratch block: bits | slot_of | sidx | sx | sy | sz | count | start (+1) | cursor
    const size_t n4 = ((size_t)n + 3) / 4 * 4, w4 = ((size_t)nwords + 3) / 4 * 4;
    const size_t words = w4 + 5 * n4 + 3 * (size_t)GREEDY_CELLS + 4;
    uint32_t *blk = nullptr;
    if (scratch_alloc((void **)&blk, words * 4, s)) return 1;
    uint32_t *bits = blk, *slot_of = bits + w4;
    int *sidx = (int *)(slot_of + n4);
    float *sx = (float *)(sidx + n4), *sy = sx + n4, *sz = sy + n4;
    int *count = (int *)(sz + n4), *start = count + GREEDY_CELLS, *cursor = start + GREEDY_CELLS + 4;
    // cells a little wider than the coverage radius: a vertex closer than the radius lies in one of the 27 cells around
    double radius = eigen_order ? sqrt((double)coverage2) : coverage;
    if (!(radius > 1e-9) || !std::isfinite(radius)) radius = 1e-9;
    const double inv_cell = 1.0 / (1.01 * radius);
    const int grid = graph_grid_for(n, 256, 4);
    cudaError_t e = cudaMemsetAsync(count, 0, (size_t)GREEDY_CELLS * 4, s);
    if (e == cudaSuccess) {
        greedy_prefilter_kernel<<<grid, 256, 0, s>>>(verts, cand, n, nodes, M, coverage, mask, bits);
        cell_count_kernel<<<grid, 256, 0, s>>>(verts, cand, n, bits, inv_cell, slot_of, count);
        cell_scan_kernel<<<1, 1024, 0, s>>>(count, start, cursor);
        cell_scatter_kernel<<<grid, 256, 0, s>>>(verts, cand, n, bits, slot_of, cursor, sidx, sx, sy, sz);
        if (eigen_order)
            greedy_select_kernel<true><<<1, GREEDY_THREADS, 0, s>>>(verts, cand, n, bits, nwords, coverage, coverage2, inv_cell, start,
                                                                   sidx, sx, sy, sz, accepted, accepted_pos, n_accepted);
        else
            greedy_select_kernel<false><<<1, GREEDY_THREADS, 0, s>>>(verts, cand, n, bits, nwords, coverage, coverage2, inv_cell, start,
                                                                    sidx, sx, sy, sz, accepted, accepted_pos, n_accepted);
        e = cudaGetLastError();
    }
    cudaFreeAsync(blk, s);
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
    nearest_node_kernel<false><<<graph_grid_for(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(points, n, nodes, M, dist, idx);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_argmin_norm(const float *queries, int64_t n, const float *points, int32_t M, float *dist, int32_t *idx,
                               void *stream)
{
    OFX_REQUIRE(n >= 0 && M >= 1, "ofx_argmin_norm: need n >= 0 and M >= 1 (n=%lld, M=%d)", (long long)n, M);
    if (n == 0) return 0;
    OFX_REQUIRE(queries && points && (dist || idx), "ofx_argmin_norm: null pointer");
    if (n <= 16384 && M >= 2048)
        argmin_norm_wide_kernel<<<(int)n, 256, 0, (cudaStream_t)stream>>>(queries, points, M, dist, idx);
    else
        nearest_node_kernel<true><<<graph_grid_for(n, 64, 16), 64, 0, (cudaStream_t)stream>>>(queries, n, points, M, dist, idx);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_pixel_anchors_euclidean(const float *nodes, int32_t M, const float *point_image, int32_t H, int32_t W,
                                           float node_coverage, int32_t *anchors, float *weights, void *stream)
{
    OFX_REQUIRE(M >= 0 && H >= 0 && W >= 0, "ofx_pixel_anchors_euclidean: bad sizes");
    const int64_t npix = (int64_t)H * W;
    if (npix == 0) return 0;
    OFX_REQUIRE(point_image && anchors && weights && (M == 0 || nodes), "ofx_pixel_anchors_euclidean: null pointer");
    pixel_anchors_kernel<<<graph_grid_for(npix, 256, 4), 256, 0, (cudaStream_t)stream>>>(nodes, M, point_image, npix, node_coverage,
                                                                                        anchors, weights);
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
    if (scratch_alloc((void **)&alive, (size_t)F, s)) return 1;
    if (scratch_alloc((void **)&cnt, (size_t)V * sizeof(int), s)) { cudaFreeAsync(alive, s); return 1; }
    cudaError_t e;
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
    const bool vec = Dz % 4 == 0 && Dz >= 32 && (int64_t)Dx * Dy < INT_MAX && (((uintptr_t)tsdf & 15) == 0) && (((uintptr_t)region & 3) == 0);
    const TruncBounds b{float_at_most(0.9), float_at_most(max_diff)};
    if (vec) truncated_region_vec_kernel<<<graph_grid_for((int64_t)Dx * Dy * 32, 256, OFX_TR_CTAS), 256, 0, s>>>(tsdf, Dx, Dy, Dz, b, region);
    else truncated_region_scalar_kernel<<<graph_grid_for(n, 256, 8), 256, 0, s>>>(tsdf, Dx, Dy, Dz, b, region);
    OFX_LAUNCH_CHECK();
    return 0;
}
