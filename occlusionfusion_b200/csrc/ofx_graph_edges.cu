// ofx_graph_edges.cu -- the edge half of the graph builder on the device:
//   ofx_compute_edges_geodesic   compute_edges_geodesic     csrc/cpu/graph_proc.cpp:155-300  (binding csrc/main.cpp:30)
//   ofx_node_and_edge_clean_up   node_and_edge_clean_up     csrc/cpu/graph_proc.cpp:388-438  (main.cpp:32)
//   ofx_compute_clusters         compute_clusters           csrc/cpu/graph_proc.cpp:440-481  (main.cpp:33)
//   ofx_upsample_index           create_graph_pyramid, up_sample_idx   fusion_with_occlusion/embedded_deformation_graph.py:279-298
//
// Geodesic edges.  The reference runs, per node, a Dijkstra search over the mesh with a std::priority_queue and stops at
// the nMaxNeighbors-th other node it pops.  What that search RETURNS depends on the heap only through ties:
//   * the distance at which a vertex is popped first is d(v) = min over mesh neighbours u of fl(d(u) + |uv|), the unique
//     fixed point of the float32 relaxation (fl is monotone, edge lengths are positive), whatever the pop order;
//   * the neighbour nodes come out in ascending d, so they are the nMaxNeighbors smallest d among the node vertices --
//     unless two of those distances are EQUAL, when the heap's internal order decides.
// So each node gets one CTA that relaxes distances in parallel (label-correcting, frontier lists, float atomicMin on the
// bit pattern) inside a radius that grows until enough nodes lie inside it, ranks the node vertices it found, and flags
// itself when a tie touches the answer; flagged nodes (regular grids; never on a real mesh) are redone by
// geodesic_exact_kernel, one lane replaying libstdc++'s push_heap / pop_heap step by step.  Edge lengths are Eigen's
// unrolled norm sqrtf(dx^2 + (dy^2 + dz^2)), sums float32, --fmad=false: the bits the reference's C++ produces.
// The weights exp(-d^2 / (2 c^2)) are normalised on the host from the distances (ofx_edge_weights_host): they go through
// the C library's expf, as the reference's do.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include "../../include/ofx.h"
#include "ofx_common.h"
#include "ofx_math.cuh"
#include "ofx_scan.cuh"

namespace ofx {

int sm_count();                                                   // ofx_points.cu
int graph_scratch_alloc(void **p, size_t bytes, cudaStream_t s);  // ofx_graph.cu: the stream-ordered scratch pool

constexpr int INF_BITS = 0x7f800000;

// ---- vertex adjacency (CSR) from the faces: sorted, unique, no self loops                       graph_proc.cpp:176-189 ----
__global__ void adj_degree_kernel(const int *__restrict__ faces, int F, int V, int *__restrict__ deg, int *__restrict__ err)
{
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const int v = faces[3 * f + j];
        if ((unsigned)v >= (unsigned)V) { atomicExch(err, -2); return; }
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) atomicAdd(deg + faces[3 * f + j], 2);
}

__global__ void adj_fill_kernel(const int *__restrict__ faces, int F, int V, const int *__restrict__ off, int *__restrict__ fill,
                                int *__restrict__ raw)
{
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    const int a = faces[3 * f], b = faces[3 * f + 1], c = faces[3 * f + 2];
    if ((unsigned)a >= (unsigned)V || (unsigned)b >= (unsigned)V || (unsigned)c >= (unsigned)V) return;   // reported by adj_degree_kernel
    int p;
    p = atomicAdd(fill + a, 2); raw[off[a] + p] = b; raw[off[a] + p + 1] = c;
    p = atomicAdd(fill + b, 2); raw[off[b] + p] = a; raw[off[b] + p + 1] = c;
    p = atomicAdd(fill + c, 2); raw[off[c] + p] = a; raw[off[c] + p + 1] = b;
}

// per vertex: sort its (short) raw list ascending, drop duplicates and itself; cnt[v] = entries left (in place)
__global__ void adj_sort_unique_kernel(int V, const int *__restrict__ off, int *__restrict__ raw, int *__restrict__ cnt)
{
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    int *a = raw + off[v];
    const int n = off[v + 1] - off[v];
    for (int i = 1; i < n; ++i) {
        const int x = a[i];
        int j = i - 1;
        while (j >= 0 && a[j] > x) { a[j + 1] = a[j]; --j; }
        a[j + 1] = x;
    }
    int m = 0;
    for (int i = 0; i < n; ++i)
        if (a[i] != v && (m == 0 || a[m - 1] != a[i])) a[m++] = a[i];
    cnt[v] = m;
}

__global__ void vertex_node_kernel(const int *__restrict__ node_idx, int first, int count, int *__restrict__ vertex_node)
{
    // mapVertexToNode (graph_proc.cpp:192-199): the LAST node that names a vertex wins -> atomicMax of the node id
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < count && node_idx[first + n] >= 0) atomicMax(vertex_node + node_idx[first + n], n);
}

__device__ __forceinline__ float edge_length(const float *__restrict__ verts, int a, int b)
{
    const float dx = verts[3 * a] - verts[3 * b], dy = verts[3 * a + 1] - verts[3 * b + 1], dz = verts[3 * a + 2] - verts[3 * b + 2];
    return sqrtf((dx * dx) + ((dy * dy) + (dz * dz)));          // Eigen's unrolled (a - b).norm()
}

// One launch serves several searches over the same mesh (the pyramid levels of create_graph_pyramid: different node
// sets, neighbour counts and radii): a job = a run of consecutive entries of node_idx with its own parameters.
constexpr int GEO_MAX_JOBS = 8;
struct GeoJob {
    int first, count;      // nodes [first, first + count) of the concatenated node list
    int E;                 // neighbours wanted
    int only_valid, enforce_total;
    float coverage, max_influence;
    int out_off;           // offset of the job's (count, E) block in edges / dists
};
struct GeoArgs {
    const float *verts;
    const uint8_t *valid;
    const int *off, *cnt, *adj, *vertex_node, *node_idx;   // vertex_node: one map of V entries per job
    int V, M, n_jobs;
    GeoJob jobs[GEO_MAX_JOBS];
    int *edges;            // per job (count, E), pre-filled by the caller
    float *dists;          // same shape
    int *n_found;          // (M,)
    uint8_t *need_exact;   // (M,) 1 = a tie touches the answer: redo with geodesic_exact_kernel
    int *error;            // [0]: vertex reached although not valid (+1), 0 = none
    // per-CTA scratch, V entries each
    int *dist_bits;        // initialised to INF_BITS, restored after every node
    int *lists;            // 4 lists: touched, frontier A, frontier B, deferred
    unsigned *flags;       // per vertex two bits, 16 vertices per word: queued in a frontier list / in the deferred list
};

constexpr int GEO_THREADS = 256;
constexpr int GEO_CAND_CAP = 1024;

__device__ __forceinline__ int job_of(const GeoArgs &g, int node)
{
    int j = 0;
    while (j + 1 < g.n_jobs && node >= g.jobs[j + 1].first) ++j;
    return j;
}

__device__ __forceinline__ unsigned flag_test_and_set(unsigned *w, int v, unsigned bit)
{
    const unsigned m = bit << (2 * (v & 15));
    return atomicOr(w + (v >> 4), m) & m;
}
__device__ __forceinline__ void flag_clear(unsigned *w, int v, unsigned bit) { atomicAnd(w + (v >> 4), ~(bit << (2 * (v & 15)))); }

// SMEM: the per-vertex distances and flags of the CTA's search live in shared memory (meshes up to ~52 k vertices: the
// relaxation is a chain of dependent atomics, 10x faster there than through L2); else in global scratch.
template <bool SMEM>
__global__ void __launch_bounds__(GEO_THREADS) geodesic_relax_kernel(GeoArgs g)
{
    extern __shared__ int s_geo[];
    __shared__ int s_n_touched, s_n_front, s_n_next, s_n_def, s_n_cand, s_tie;
    __shared__ float s_cd[GEO_CAND_CAP];
    __shared__ int s_cn[GEO_CAND_CAP];
    const int n_words = (g.V + 15) >> 4;
    int *dist = SMEM ? s_geo : g.dist_bits + (size_t)blockIdx.x * g.V;
    unsigned *flag = SMEM ? reinterpret_cast<unsigned *>(s_geo + g.V) : g.flags + (size_t)blockIdx.x * n_words;
    int *touched = g.lists + (size_t)blockIdx.x * 4 * g.V;
    int *front = touched + g.V, *next = front + g.V, *deferred = next + g.V;
    const int tid = threadIdx.x;
    if (SMEM) {
        for (int i = tid; i < g.V; i += GEO_THREADS) dist[i] = INF_BITS;
        for (int i = tid; i < n_words; i += GEO_THREADS) flag[i] = 0u;
        __syncthreads();
    }

    for (int node = blockIdx.x; node < g.M; node += gridDim.x) {
        const int v0 = g.node_idx[node];
        const int jb = job_of(g, node);
        const GeoJob job = g.jobs[jb];
        const int self = node - job.first;                        // the node's id within its job
        const int *vnode = g.vertex_node + (size_t)jb * g.V;
        if (tid == 0) { g.need_exact[node] = 0; g.n_found[node] = 0; }
        if (v0 < 0) continue;                                                                 // graph_proc.cpp:215
        if (job.only_valid && !g.valid[v0]) { if (tid == 0) atomicExch(g.error, v0 + 1); continue; }   // the reference exit(0)s, :233-236
        if (tid == 0) {
            dist[v0] = 0;
            touched[0] = v0; front[0] = v0;
            s_n_touched = 1; s_n_front = 1; s_n_next = 0; s_n_def = 0;
        }
        float T = job.enforce_total ? 2.0f * job.coverage : job.max_influence;
        __syncthreads();
        for (;;) {
            // ---- relax inside the radius T until nothing changes ----
            while (s_n_front > 0) {
                const int nf = s_n_front;
                for (int i = tid; i < nf; i += GEO_THREADS) {
                    const int u = front[i];
                    flag_clear(flag, u, 1u);                                  // leaves the queue BEFORE its distance is read: a later
                    __threadfence_block();                                    // improvement of u queues it again
                    const float du = __int_as_float(*reinterpret_cast<volatile int *>(dist + u));
                    if (du > T) {
                        if (!flag_test_and_set(flag, u, 2u)) deferred[atomicAdd(&s_n_def, 1)] = u;
                        continue;
                    }
                    const int o = g.off[u], c = g.cnt[u];
                    for (int k = 0; k < c; ++k) {
                        const int nb = g.adj[o + k];
                        if (job.only_valid && !g.valid[nb]) continue;                         // :258-261
                        const float nd = du + edge_length(g.verts, u, nb);                    // :264
                        if (!job.enforce_total && nd > job.max_influence) continue;           // :266-274
                        const int nbits = __float_as_int(nd);
                        const int old = atomicMin(dist + nb, nbits);
                        if (nbits < old) {
                            if (old == INF_BITS) touched[atomicAdd(&s_n_touched, 1)] = nb;
                            if (!flag_test_and_set(flag, nb, 1u)) next[atomicAdd(&s_n_next, 1)] = nb;
                        }
                    }
                }
                __syncthreads();
                if (tid == 0) { s_n_front = s_n_next; s_n_next = 0; }
                int *t = front; front = next; next = t;
                __syncthreads();
            }
            // ---- the node vertices inside T ----
            if (tid == 0) { s_n_cand = 0; s_tie = 0; }
            __syncthreads();
            const int nt = s_n_touched;
            for (int i = tid; i < nt; i += GEO_THREADS) {
                const int v = touched[i];
                const int other = vnode[v];
                const float d = __int_as_float(dist[v]);
                if (other >= 0 && other != self && d <= T) {                                  // :239
                    const int j = atomicAdd(&s_n_cand, 1);
                    if (j < GEO_CAND_CAP) { s_cd[j] = d; s_cn[j] = other; }
                }
            }
            __syncthreads();
            const bool done = !job.enforce_total || s_n_cand >= job.E || s_n_def == 0;
            if (done) break;
            // not enough nodes yet: widen the radius; the vertices that waited outside it go back to work
            T *= 1.7f;
            const int nd = s_n_def;
            for (int i = tid; i < nd; i += GEO_THREADS) {
                const int u = deferred[i];
                front[i] = u;
                flag_clear(flag, u, 2u);                                      // no longer deferred ...
                flag_test_and_set(flag, u, 1u);                               // ... queued again
            }
            __syncthreads();
            if (tid == 0) { s_n_front = nd; s_n_def = 0; }
            __syncthreads();
        }
        // ---- rank the candidates: the E nearest, ascending; equal distances that touch the answer -> exact replay ----
        const int nc = s_n_cand;
        if (nc > GEO_CAND_CAP) {
            if (tid == 0) g.need_exact[node] = 1;
        } else {
            for (int i = tid; i < nc; i += GEO_THREADS) {
                const float d = s_cd[i];
                int rank = 0, ties = 0;
                for (int j = 0; j < nc; ++j) {
                    const float e = s_cd[j];
                    rank += (e < d) || (e == d && j < i);
                    ties += (e == d && j != i);
                }
                // this candidate's rank without its ties is rank - (ties before it): a tie group that starts inside
                // the first E entries changes the answer's order or membership
                if (ties) {
                    int first = 0;
                    for (int j = 0; j < nc; ++j) first += s_cd[j] < d;
                    if (first < job.E) s_tie = 1;
                }
                if (rank < job.E) {
                    g.edges[job.out_off + (size_t)self * job.E + rank] = s_cn[i];
                    g.dists[job.out_off + (size_t)self * job.E + rank] = d;
                }
            }
            __syncthreads();
            if (tid == 0) { g.n_found[node] = min(nc, job.E); g.need_exact[node] = (uint8_t)s_tie; }
        }
        // ---- restore the scratch ----
        __syncthreads();
        const int nt = s_n_touched;
        for (int i = tid; i < nt; i += GEO_THREADS) {
            const int v = touched[i];
            dist[v] = INF_BITS;
            flag_clear(flag, v, 3u);
        }
        __syncthreads();
    }
}

// The reference's search replayed step by step for the nodes the relaxation flagged (ties): one warp per node, lane 0
// walks; the heap is libstdc++'s (push_back + __push_heap; __pop_heap = __adjust_heap + __push_heap) on (vertex, distance)
// with CustomCompare (left.second > right.second), the visited set is a stamp per vertex.
struct HeapItem { int v; float d; };

__device__ __forceinline__ void heap_push(HeapItem *h, int &n, HeapItem x)
{
    int hole = n++;
    while (hole > 0) {
        const int parent = (hole - 1) / 2;
        if (!(h[parent].d > x.d)) break;
        h[hole] = h[parent];
        hole = parent;
    }
    h[hole] = x;
}

__device__ __forceinline__ HeapItem heap_pop(HeapItem *h, int &n)
{
    const HeapItem top = h[0];
    const int len = --n;                         // elements left; the old last one is re-inserted from the root
    if (len > 0) {
        const HeapItem value = h[len];
        int hole = 0, child = 0;
        while (child < (len - 1) / 2) {
            child = 2 * (child + 1);
            if (h[child].d > h[child - 1].d) --child;
            h[hole] = h[child];
            hole = child;
        }
        if ((len & 1) == 0 && child == (len - 2) / 2) {
            child = 2 * (child + 1);
            h[hole] = h[child - 1];
            hole = child - 1;
        }
        while (hole > 0) {                       // __push_heap(first, hole, 0, value)
            const int parent = (hole - 1) / 2;
            if (!(h[parent].d > value.d)) break;
            h[hole] = h[parent];
            hole = parent;
        }
        h[hole] = value;
    }
    return top;
}

__global__ void __launch_bounds__(32) geodesic_exact_kernel(GeoArgs g, const int *__restrict__ todo, int n_todo, int *__restrict__ stamps,
                                                            HeapItem *__restrict__ heaps, int heap_cap)
{
    if (threadIdx.x != 0) return;
    int *stamp = stamps + (size_t)blockIdx.x * g.V;
    HeapItem *h = heaps + (size_t)blockIdx.x * heap_cap;
    for (int t = blockIdx.x; t < n_todo; t += gridDim.x) {
        const int node = todo[t];
        const int v0 = g.node_idx[node];
        const int jb = job_of(g, node);
        const GeoJob job = g.jobs[jb];
        const int self = node - job.first;
        const int *vnode = g.vertex_node + (size_t)jb * g.V;
        int n = 0, cnt = 0;
        heap_push(h, n, HeapItem{v0, 0.f});
        while (n > 0) {
            const HeapItem cur = heap_pop(h, n);
            if (stamp[cur.v] == node + 1) continue;                                           // :231 (stamps start at 0)
            if (job.only_valid && !g.valid[cur.v]) { atomicExch(g.error, cur.v + 1); break; }
            const int other = vnode[cur.v];
            if (other >= 0 && other != self) {                                                // :239-245
                g.edges[job.out_off + (size_t)self * job.E + cnt] = other;
                g.dists[job.out_off + (size_t)self * job.E + cnt] = cur.d;
                if (++cnt >= job.E) break;
            }
            stamp[cur.v] = node + 1;
            const int o = g.off[cur.v], c = g.cnt[cur.v];
            for (int k = 0; k < c; ++k) {
                const int nb = g.adj[o + k];
                if (job.only_valid && !g.valid[nb]) continue;
                const float d = cur.d + edge_length(g.verts, cur.v, nb);
                if (job.enforce_total || d <= job.max_influence) {
                    if (n >= heap_cap) { atomicExch(g.error, -1); n = 0; cnt = 0; break; }     // cannot happen: capacity = all mesh edges + 1
                    heap_push(h, n, HeapItem{nb, d});
                }
            }
        }
        g.n_found[node] = cnt;
    }
}

__global__ void fill_int_kernel(int *p, int64_t n, int v)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

__global__ void compact_flags_kernel(const uint8_t *__restrict__ flag, int M, int *__restrict__ list, int *__restrict__ n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < M && flag[i]) list[atomicAdd(n, 1)] = i;
}

// ---- node_and_edge_clean_up (graph_proc.cpp:388-438): drop nodes left with at most one neighbour, repeatedly.  The
// reference sweeps in node order with immediate effect; removal only ever lowers other nodes' counts, so the set left
// at the fixed point does not depend on the order and parallel sweeps reach the same one.  One CTA.
__global__ void __launch_bounds__(1024) clean_up_kernel(const int *__restrict__ edges, int M, int E, uint8_t *__restrict__ valid,
                                                        uint8_t *__restrict__ removed)
{
    __shared__ int s_changed;
    for (int n = threadIdx.x; n < M; n += blockDim.x) removed[n] = 0;
    __syncthreads();
    for (;;) {
        if (threadIdx.x == 0) s_changed = 0;
        __syncthreads();
        for (int n = threadIdx.x; n < M; n += blockDim.x) {
            if (!valid[n]) continue;
            int cnt = 0;
            for (int i = 0; i < E; ++i) {
                const int nb = edges[(size_t)n * E + i];
                if (nb == -1) break;                                                          // :408
                if (nb >= 0 && nb < M && *reinterpret_cast<volatile uint8_t *>(removed + nb)) continue;
                ++cnt;
            }
            if (cnt <= 1) { valid[n] = 0; removed[n] = 1; s_changed = 1; }                    // :419-424
        }
        __syncthreads();
        if (!s_changed) break;
        __syncthreads();
    }
}

// ---- compute_clusters (graph_proc.cpp:440-481): connected components of the undirected edge set, numbered in the order
// of their smallest node.  Min-label propagation to the fixed point, then the roots are ranked.  One CTA.
__global__ void __launch_bounds__(1024) clusters_kernel(const int *__restrict__ edges, int M, int E, int *__restrict__ label,
                                                        int *__restrict__ clusters, int *__restrict__ sizes, int *__restrict__ n_clusters,
                                                        int *__restrict__ err)
{
    __shared__ int s_changed, s_count;
    for (int n = threadIdx.x; n < M; n += blockDim.x) { label[n] = n; sizes[n] = 0; }
    __syncthreads();
    for (;;) {
        if (threadIdx.x == 0) s_changed = 0;
        __syncthreads();
        for (int n = threadIdx.x; n < M; n += blockDim.x)
            for (int i = 0; i < E; ++i) {
                const int nb = edges[(size_t)n * E + i];
                if (nb == -1) break;                                                          // :455
                if (nb < 0 || nb >= M) { atomicExch(err, n + 1); break; }
                const int a = *reinterpret_cast<volatile int *>(label + n), b = *reinterpret_cast<volatile int *>(label + nb);
                if (a < b) { atomicMin(label + nb, a); s_changed = 1; }
                else if (b < a) { atomicMin(label + n, b); s_changed = 1; }
            }
        __syncthreads();
        // pointer jumping: every label moves to its label's label
        for (int n = threadIdx.x; n < M; n += blockDim.x) {
            int l = label[n];
            while (label[l] != l) l = label[l];
            label[n] = l;
        }
        __syncthreads();
        if (!s_changed) break;
        __syncthreads();
    }
    // roots (label[n] == n) in node order get the cluster numbers 0, 1, ...: serial prefix count by one thread per chunk
    if (threadIdx.x == 0) {
        int c = 0;
        for (int n = 0; n < M; ++n)
            if (label[n] == n) clusters[n] = c++;
        s_count = c;
        *n_clusters = c;
    }
    __syncthreads();
    for (int n = threadIdx.x; n < M; n += blockDim.x) {
        const int r = label[n];
        if (r != n) clusters[n] = clusters[r];
    }
    __syncthreads();
    for (int n = threadIdx.x; n < M; n += blockDim.x) atomicAdd(sizes + clusters[n], 1);
    (void)s_count;
}

// ---- up_sample_idx of create_graph_pyramid (embedded_deformation_graph.py:279-298): for node i the position, within
// the kept list, of the kept node nearest to it among those kept BEFORE i was examined (the first node maps to itself);
// float32 norms sqrtf((dx^2+dy^2)+dz^2), first index among equal norms.
__global__ void upsample_index_kernel(const float *__restrict__ nodes, int M, const int *__restrict__ down_idx, int n_down,
                                      int *__restrict__ up_idx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    const float px = nodes[3 * i], py = nodes[3 * i + 1], pz = nodes[3 * i + 2];
    float best = INFINITY;
    int bi = 0;
    for (int k = 0; k < n_down; ++k) {
        const int q = down_idx[k];
        if (q >= i) break;                                           // kept nodes come in node order: only those before i
        const float dx = nodes[3 * q] - px, dy = nodes[3 * q + 1] - py, dz = nodes[3 * q + 2] - pz;
        const float d = sqrtf(((dx * dx) + (dy * dy)) + (dz * dz));
        if (d < best) { best = d; bi = k; }
    }
    up_idx[i] = bi;
}

}  // namespace ofx

using namespace ofx;

extern "C" int ofx_compute_edges_geodesic_batch(const float *verts, const uint8_t *valid, int32_t V, const int32_t *faces, int32_t F,
                                                const int32_t *node_idx, int32_t n_jobs, const int32_t *job_count_host,
                                                const int32_t *job_neighbours_host, const float *job_coverage_host,
                                                const int32_t *job_flags_host, int32_t *edges, float *distances, int32_t *n_found,
                                                int32_t *status_host, void *stream)
{
    OFX_REQUIRE(V >= 0 && F >= 0 && n_jobs >= 1 && n_jobs <= GEO_MAX_JOBS, "ofx_compute_edges_geodesic: bad sizes (1..%d jobs)", GEO_MAX_JOBS);
    OFX_REQUIRE(job_count_host && job_neighbours_host && job_coverage_host && job_flags_host, "ofx_compute_edges_geodesic: null job array");
    if (status_host) status_host[0] = status_host[1] = 0;
    GeoArgs g{};
    int M = 0, out = 0, any_only_valid = 0;
    for (int j = 0; j < n_jobs; ++j) {
        OFX_REQUIRE(job_count_host[j] >= 0 && job_neighbours_host[j] >= 1, "ofx_compute_edges_geodesic: bad job %d", j);
        g.jobs[j] = GeoJob{M, job_count_host[j], job_neighbours_host[j], job_flags_host[j] & 1, (job_flags_host[j] >> 1) & 1,
                           job_coverage_host[j], 2.f * job_coverage_host[j], out};                   // :173
        any_only_valid |= job_flags_host[j] & 1;
        M += job_count_host[j];
        out += job_count_host[j] * job_neighbours_host[j];
    }
    if (M == 0) return 0;
    OFX_REQUIRE(verts && faces && node_idx && edges && distances && n_found && (!any_only_valid || valid) && V > 0,
                "ofx_compute_edges_geodesic: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    // distances + flags of a search in shared memory when the mesh fits (4.25 bytes per vertex), else in global scratch
    const int n_words = (V + 15) / 16;
    const size_t geo_smem = ((size_t)V + n_words) * sizeof(int);
    const bool in_smem = geo_smem <= 200 * 1024;
    const int per_sm = in_smem ? std::max(1, std::min(4, (int)((220 * 1024) / (geo_smem + 10 * 1024)))) : 2;
    const int grid = std::min(M, per_sm * sm_count());
    // scratch: adjacency + per-CTA search state
    const size_t n_raw = (size_t)6 * F + 4;
    const int n_scan_sums = (V + 1 + SCAN_BLOCK * SCAN_ITEMS - 1) / (SCAN_BLOCK * SCAN_ITEMS) + 2;
    int *deg = nullptr, *off = nullptr, *cnt = nullptr, *raw = nullptr, *sums = nullptr, *vnode = nullptr, *small = nullptr;
    int *dist_bits = nullptr, *lists = nullptr, *todo = nullptr;
    unsigned *flags = nullptr;
    uint8_t *need = nullptr;
    if (graph_scratch_alloc((void **)&deg, ((size_t)V + 2) * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&off, ((size_t)V + 2) * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&cnt, ((size_t)V + 2) * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&raw, n_raw * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&sums, (size_t)n_scan_sums * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&vnode, (size_t)n_jobs * V * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&small, 8 * sizeof(int), s)) return 1;        // [0] error, [1] number of flagged nodes, [2] total degree
    if (graph_scratch_alloc((void **)&dist_bits, in_smem ? 16 : (size_t)grid * V * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&lists, (size_t)grid * 4 * V * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&flags, in_smem ? 16 : (size_t)grid * n_words * sizeof(unsigned), s)) return 1;
    if (graph_scratch_alloc((void **)&need, (size_t)M + 4, s)) return 1;
    if (graph_scratch_alloc((void **)&todo, ((size_t)M + 1) * sizeof(int), s)) return 1;
    OFX_CUDA(cudaMemsetAsync(deg, 0, ((size_t)V + 2) * sizeof(int), s));
    OFX_CUDA(cudaMemsetAsync(small, 0, 8 * sizeof(int), s));
    if (!in_smem) OFX_CUDA(cudaMemsetAsync(flags, 0, (size_t)grid * n_words * sizeof(unsigned), s));
    if (F > 0) adj_degree_kernel<<<(F + 255) / 256, 256, 0, s>>>(faces, F, V, deg, small);
    if (int rc = exclusive_scan(deg, V + 1, off, sums, small + 2, s)) return rc;
    OFX_CUDA(cudaMemsetAsync(cnt, 0, ((size_t)V + 2) * sizeof(int), s));
    if (F > 0) adj_fill_kernel<<<(F + 255) / 256, 256, 0, s>>>(faces, F, V, off, cnt, raw);
    adj_sort_unique_kernel<<<(V + 127) / 128, 128, 0, s>>>(V, off, raw, cnt);
    fill_int_kernel<<<std::min((n_jobs * V + 255) / 256, 4 * sm_count()), 256, 0, s>>>(vnode, (int64_t)n_jobs * V, -1);
    for (int j = 0; j < n_jobs; ++j)
        if (g.jobs[j].count > 0)
            vertex_node_kernel<<<(g.jobs[j].count + 255) / 256, 256, 0, s>>>(node_idx, g.jobs[j].first, g.jobs[j].count, vnode + (size_t)j * V);
    if (!in_smem) fill_int_kernel<<<4 * sm_count(), 256, 0, s>>>(dist_bits, (int64_t)grid * V, INF_BITS);
    OFX_LAUNCH_CHECK();

    g.verts = verts; g.valid = valid; g.off = off; g.cnt = cnt; g.adj = raw; g.vertex_node = vnode; g.node_idx = node_idx;
    g.V = V; g.M = M; g.n_jobs = n_jobs;
    g.edges = edges; g.dists = distances; g.n_found = n_found; g.need_exact = need; g.error = small;
    g.dist_bits = dist_bits; g.lists = lists; g.flags = flags;
    if (in_smem) {
        OFX_CUDA(cudaFuncSetAttribute(geodesic_relax_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)geo_smem));
        geodesic_relax_kernel<true><<<grid, GEO_THREADS, geo_smem, s>>>(g);
    } else {
        geodesic_relax_kernel<false><<<grid, GEO_THREADS, 0, s>>>(g);
    }
    compact_flags_kernel<<<(M + 255) / 256, 256, 0, s>>>(need, M, todo, small + 1);
    OFX_LAUNCH_CHECK();
    int host[4] = {0, 0, 0, 0};
    OFX_CUDA(cudaMemcpyAsync(host, small, 3 * sizeof(int), cudaMemcpyDeviceToHost, s));
    OFX_CUDA(cudaStreamSynchronize(s));
    int rc = 0;
    if (host[0] == 0 && host[1] > 0) {
        // ties: replay those nodes exactly (one warp each); stamps + heaps (capacity: every mesh edge once + the start)
        const int n_todo = host[1], heap_cap = host[2] + 2;
        const int egrid = std::min(n_todo, 4 * sm_count());
        int *stamps = nullptr;
        HeapItem *heaps = nullptr;
        if (graph_scratch_alloc((void **)&stamps, (size_t)egrid * V * sizeof(int), s)) return 1;
        if (graph_scratch_alloc((void **)&heaps, (size_t)egrid * heap_cap * sizeof(HeapItem), s)) return 1;
        OFX_CUDA(cudaMemsetAsync(stamps, 0, (size_t)egrid * V * sizeof(int), s));
        geodesic_exact_kernel<<<egrid, 32, 0, s>>>(g, todo, n_todo, stamps, heaps, heap_cap);
        OFX_LAUNCH_CHECK();
        OFX_CUDA(cudaMemcpyAsync(host, small, sizeof(int), cudaMemcpyDeviceToHost, s));
        OFX_CUDA(cudaStreamSynchronize(s));
        cudaFreeAsync(stamps, s); cudaFreeAsync(heaps, s);
    }
    if (status_host) { status_host[0] = host[0]; status_host[1] = host[1]; }
    for (void *p : {(void *)deg, (void *)off, (void *)cnt, (void *)raw, (void *)sums, (void *)vnode, (void *)small, (void *)dist_bits,
                    (void *)lists, (void *)flags, (void *)need, (void *)todo})
        cudaFreeAsync(p, s);
    if (host[0] > 0)
        rc = fail("ofx_compute_edges_geodesic: reached vertex %d, which is not a valid vertex", host[0] - 1);
    else if (host[0] < 0)
        rc = fail("ofx_compute_edges_geodesic: a face index is out of range or the search heap overflowed (%d)", host[0]);
    return rc;
}

extern "C" int ofx_compute_edges_geodesic(const float *verts, const uint8_t *valid, int32_t V, const int32_t *faces, int32_t F,
                                          const int32_t *node_idx, int32_t M, int32_t max_neighbours, float node_coverage,
                                          int32_t *edges, float *distances, int32_t *n_found, int32_t only_valid,
                                          int32_t enforce_total, int32_t *status_host, void *stream)
{
    OFX_REQUIRE(M >= 0 && max_neighbours >= 1, "ofx_compute_edges_geodesic: bad sizes");
    const int32_t flags = (only_valid ? 1 : 0) | (enforce_total ? 2 : 0);
    return ofx_compute_edges_geodesic_batch(verts, valid, V, faces, F, node_idx, 1, &M, &max_neighbours, &node_coverage, &flags, edges,
                                            distances, n_found, status_host, stream);
}

// weights of the edges from their geodesic distances, normalised per node (graph_proc.cpp:150-152, 281-296): HOST arrays,
// the C library's expf as in the reference
extern "C" int ofx_edge_weights_host(const float *distances_host, const int32_t *n_found_host, int32_t M, int32_t E,
                                     float node_coverage, float *weights_host)
{
    OFX_REQUIRE(M >= 0 && E >= 1, "ofx_edge_weights_host: bad sizes");
    if (M == 0) return 0;
    OFX_REQUIRE(distances_host && n_found_host && weights_host, "ofx_edge_weights_host: null pointer");
    for (int n = 0; n < M; ++n) {
        const int cnt = std::min(std::max(n_found_host[n], 0), E);
        float w[OFX_MAX_K * 4];
        float sum = 0.f;
        for (int i = 0; i < cnt; ++i) {
            const float d = distances_host[(size_t)n * E + i];
            const float wi = std::exp(-(d * d) / (2.f * node_coverage * node_coverage));
            if (i < (int)(sizeof(w) / sizeof(w[0]))) w[i] = wi;
            sum += wi;
        }
        OFX_REQUIRE(cnt <= (int)(sizeof(w) / sizeof(w[0])), "ofx_edge_weights_host: more than %d neighbours", (int)(sizeof(w) / sizeof(w[0])));
        if (sum > 0)
            for (int i = 0; i < cnt; ++i) weights_host[(size_t)n * E + i] = w[i] / sum;
        else if (cnt > 0)
            for (int i = 0; i < cnt; ++i) weights_host[(size_t)n * E + i] = w[i] / cnt;
    }
    return 0;
}

extern "C" int ofx_node_and_edge_clean_up(const int32_t *edges, int32_t M, int32_t E, uint8_t *valid_mask, void *stream)
{
    OFX_REQUIRE(M >= 0 && E >= 0, "ofx_node_and_edge_clean_up: bad sizes");
    if (M == 0) return 0;
    OFX_REQUIRE(edges && valid_mask, "ofx_node_and_edge_clean_up: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    uint8_t *removed = nullptr;
    if (graph_scratch_alloc((void **)&removed, (size_t)M + 4, s)) return 1;
    clean_up_kernel<<<1, 1024, 0, s>>>(edges, M, E, valid_mask, removed);
    OFX_LAUNCH_CHECK();
    cudaFreeAsync(removed, s);
    return 0;
}

extern "C" int ofx_compute_clusters(const int32_t *edges, int32_t M, int32_t E, int32_t *clusters, int32_t *sizes,
                                    int32_t *n_clusters, void *stream)
{
    OFX_REQUIRE(M >= 0 && E >= 0 && n_clusters, "ofx_compute_clusters: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    if (M == 0) { OFX_CUDA(cudaMemsetAsync(n_clusters, 0, sizeof(int), s)); return 0; }
    OFX_REQUIRE(edges && clusters && sizes, "ofx_compute_clusters: null pointer");
    int *label = nullptr, *err = nullptr;
    if (graph_scratch_alloc((void **)&label, (size_t)M * sizeof(int), s)) return 1;
    if (graph_scratch_alloc((void **)&err, sizeof(int), s)) return 1;
    OFX_CUDA(cudaMemsetAsync(err, 0, sizeof(int), s));
    clusters_kernel<<<1, 1024, 0, s>>>(edges, M, E, label, clusters, sizes, n_clusters, err);
    OFX_LAUNCH_CHECK();
    int e = 0;
    OFX_CUDA(cudaMemcpyAsync(&e, err, sizeof(int), cudaMemcpyDeviceToHost, s));
    OFX_CUDA(cudaStreamSynchronize(s));
    cudaFreeAsync(label, s); cudaFreeAsync(err, s);
    OFX_REQUIRE(e == 0, "ofx_compute_clusters: an edge of node %d is out of range", e - 1);
    return 0;
}

extern "C" int ofx_upsample_index(const float *nodes, int32_t M, const int32_t *down_idx, int32_t n_down, int32_t *up_idx,
                                  void *stream)
{
    OFX_REQUIRE(M >= 0 && n_down >= 0, "ofx_upsample_index: bad sizes");
    if (M == 0) return 0;
    OFX_REQUIRE(nodes && down_idx && up_idx, "ofx_upsample_index: null pointer");
    upsample_index_kernel<<<(M + 127) / 128, 128, 0, (cudaStream_t)stream>>>(nodes, M, down_idx, n_down, up_idx);
    OFX_LAUNCH_CHECK();
    return 0;
}

// ---- ofx_compute_clusters_noalloc: the same kernel with its label scratch in a file-scope device array ----------------
// No counterpart in the reference; added for callers that cannot allocate (CUDA
// graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the array
// (OFX_NOALLOC_MAX_NODES ints = 256 KiB) until exit; every call shares it, so calls that overlap on one GPU race.
namespace ofx {
static __device__ __align__(256) int g_noalloc_labels[OFX_NOALLOC_MAX_NODES];
constexpr int NOALLOC_MAX_DEVICES = 64;
static int *g_noalloc_labels_ptr[NOALLOC_MAX_DEVICES];      // device address of the array, per device (host code cannot use the name)
}  // namespace ofx

extern "C" int ofx_compute_clusters_noalloc(const int32_t *edges, int32_t M, int32_t E, int32_t *clusters, int32_t *sizes,
                                            int32_t *n_clusters, int32_t *status, void *stream)
{
    OFX_REQUIRE(M >= 0 && E >= 0 && n_clusters && status, "ofx_compute_clusters_noalloc: bad arguments");
    OFX_REQUIRE(M <= OFX_NOALLOC_MAX_NODES, "ofx_compute_clusters_noalloc: %d nodes exceed the static scratch (%d); use ofx_compute_clusters",
                M, OFX_NOALLOC_MAX_NODES);
    cudaStream_t s = (cudaStream_t)stream;
    OFX_CUDA(cudaMemsetAsync(status, 0, sizeof(int), s));                      // the kernel only ever raises it
    if (M == 0) { OFX_CUDA(cudaMemsetAsync(n_clusters, 0, sizeof(int), s)); return 0; }
    OFX_REQUIRE(edges && clusters && sizes, "ofx_compute_clusters_noalloc: null pointer");
    int dev = 0;
    OFX_CUDA(cudaGetDevice(&dev));
    OFX_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "ofx_compute_clusters_noalloc: device %d out of range", dev);
    if (!g_noalloc_labels_ptr[dev]) {                                          // once per device, with that device current
        int *p = nullptr;
        OFX_CUDA(cudaGetSymbolAddress((void **)&p, g_noalloc_labels));
        OFX_REQUIRE(p != nullptr, "ofx_compute_clusters_noalloc: no device address for the scratch array");
        g_noalloc_labels_ptr[dev] = p;
    }
    // the kernel writes label[0..M) before it reads it: nothing of the previous call's contents is used
    clusters_kernel<<<1, 1024, 0, s>>>(edges, M, E, g_noalloc_labels_ptr[dev], clusters, sizes, n_clusters, status);
    OFX_LAUNCH_CHECK();
    return 0;
}
