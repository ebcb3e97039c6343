// ofx_graph_host.cu -- the host half of the graph builder in libofx.so: geodesic edges
// between graph nodes, the "not enough neighbours" clean-up and the connected-component clusters.  In the reference these
// are serial host C++ behind pybind11 (csrc/cpu/graph_proc.cpp:155-300, 388-438, 440-481; bindings csrc/main.cpp:30, 32, 33);
// here they are host C++ behind the C ABI (pointers are HOST pointers, names end in _host), the per-node searches of the
// geodesic edges spread over host threads.  No device work, no device needed.
//   The arithmetic that decides results follows the code it replaces: edge length = sqrtf(dx^2 + (dy^2 + dz^2)) (Eigen's
// unrolled norm of a Vector3f), path length accumulated from the node outwards in float32, neighbours of a vertex visited
// in ascending index order, a binary heap with the standard library's push_heap / pop_heap on the distance alone (so
// equal distances leave the heap in the same order), weight = exp(-(d*d) / (2.f*c*c)) in float32.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>
#include "../../include/ofx.h"
#include "ofx_common.h"

namespace {

struct HeapItem {
    int v;
    float d;
};
struct FartherFirst {
    bool operator()(const HeapItem &a, const HeapItem &b) const { return a.d > b.d; }
};

// vertex -> sorted unique neighbour vertices (CSR), from the faces                          graph_proc.cpp:176-189
void build_adjacency(const int32_t *faces, int F, int V, std::vector<int> &off, std::vector<int> &adj)
{
    std::vector<int> deg(V + 1, 0);
    for (int f = 0; f < F; ++f)
        for (int j = 0; j < 3; ++j) deg[faces[3 * f + j]] += 2;
    off.assign(V + 1, 0);
    for (int v = 0; v < V; ++v) off[v + 1] = off[v] + deg[v];
    std::vector<int> raw(off[V]), fill(off.begin(), off.end() - 1);
    for (int f = 0; f < F; ++f)
        for (int j = 0; j < 3; ++j) {
            const int v = faces[3 * f + j];
            for (int k = 0; k < 3; ++k)
                if (k != j) raw[fill[v]++] = faces[3 * f + k];
        }
    adj.clear();
    adj.reserve(raw.size() / 2 + 1);
    std::vector<int> new_off(V + 1, 0);
    for (int v = 0; v < V; ++v) {
        std::sort(raw.begin() + off[v], raw.begin() + off[v + 1]);
        int last = -1;
        bool first = true;
        for (int i = off[v]; i < off[v + 1]; ++i) {
            if (raw[i] == v) continue;                                  // v_idx == n_idx (degenerate face)  :184
            if (first || raw[i] != last) { adj.push_back(raw[i]); last = raw[i]; first = false; }
        }
        new_off[v + 1] = (int)adj.size();
    }
    off.swap(new_off);
}

inline float edge_length(const float *a, const float *b)
{
    const float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return std::sqrt((dx * dx) + ((dy * dy) + (dz * dz)));
}

}  // namespace

extern "C" int ofx_compute_edges_geodesic_host(const float *verts_host, const uint8_t *valid_host, int32_t V,
                                               const int32_t *faces_host, int32_t F, const int32_t *node_idx_host, int32_t M,
                                               int32_t max_neighbours, float node_coverage, int32_t *edges_host,
                                               float *weights_host, float *distances_host, float *node_to_vertex_host,
                                               int32_t only_valid, int32_t enforce_total, int32_t n_threads)
{
    OFX_REQUIRE(V >= 0 && F >= 0 && M >= 0 && max_neighbours >= 1, "ofx_compute_edges_geodesic_host: bad sizes");
    if (M == 0) return 0;
    OFX_REQUIRE(verts_host && faces_host && node_idx_host && edges_host && weights_host && distances_host &&
                    (!only_valid || valid_host),
                "ofx_compute_edges_geodesic_host: null pointer");
    for (int i = 0; i < 3 * F; ++i)
        OFX_REQUIRE((uint32_t)faces_host[i] < (uint32_t)V, "ofx_compute_edges_geodesic_host: face index %d out of range", faces_host[i]);
    for (int i = 0; i < M; ++i)
        OFX_REQUIRE(node_idx_host[i] < V, "ofx_compute_edges_geodesic_host: node vertex %d out of range", node_idx_host[i]);
    std::vector<int> off, adj;
    build_adjacency(faces_host, F, V, off, adj);
    std::vector<int> vertex_node(V, -1);                                                     // :192-199
    for (int n = 0; n < M; ++n)
        if (node_idx_host[n] >= 0) vertex_node[node_idx_host[n]] = n;
    const float max_influence = 2.f * node_coverage;                                         // :173
    const int E = max_neighbours;
    std::atomic<int> next(0), bad_vertex(-1);
    auto worker = [&]() {
        std::vector<int> stamp(V, -1);          // stamp[v] == node  <=>  v was visited in this node's search
        std::vector<HeapItem> heap;
        std::vector<int> ids(E);
        std::vector<float> wts(E), dst(E);
        for (int node = next.fetch_add(1); node < M; node = next.fetch_add(1)) {
            const int v0 = node_idx_host[node];
            if (v0 < 0) continue;                                                             // :215
            heap.clear();
            heap.push_back(HeapItem{v0, 0.f});
            int cnt = 0;
            while (!heap.empty()) {
                std::pop_heap(heap.begin(), heap.end(), FartherFirst());
                const HeapItem cur = heap.back();
                heap.pop_back();
                if (stamp[cur.v] == node) continue;                                           // :231
                if (only_valid && !valid_host[cur.v]) { bad_vertex.store(cur.v); return; }    // the reference exit(0)s here, :233-236
                const int other = vertex_node[cur.v];
                if (other >= 0 && other != node) {                                            // :239-245
                    ids[cnt] = other;
                    wts[cnt] = std::exp(-(cur.d * cur.d) / (2.f * node_coverage * node_coverage));
                    dst[cnt] = cur.d;
                    if (++cnt >= E) break;
                }
                if (node_to_vertex_host) node_to_vertex_host[(int64_t)node * V + cur.v] = cur.d;   // :248
                stamp[cur.v] = node;
                const float *p = verts_host + 3 * (int64_t)cur.v;
                for (int i = off[cur.v]; i < off[cur.v + 1]; ++i) {
                    const int nb = adj[i];
                    if (only_valid && !valid_host[nb]) continue;                              // :258-261
                    const float d = cur.d + edge_length(p, verts_host + 3 * (int64_t)nb);
                    if (enforce_total || d <= max_influence) {                                // :266-274
                        heap.push_back(HeapItem{nb, d});
                        std::push_heap(heap.begin(), heap.end(), FartherFirst());
                    }
                }
            }
            float sum = 0.f;
            for (int i = 0; i < cnt; ++i) { edges_host[(int64_t)node * E + i] = ids[i]; sum += wts[i]; }
            if (sum > 0)                                                                      // :286-295
                for (int i = 0; i < cnt; ++i) weights_host[(int64_t)node * E + i] = wts[i] / sum;
            else if (cnt > 0)
                for (int i = 0; i < cnt; ++i) weights_host[(int64_t)node * E + i] = wts[i] / cnt;
            for (int i = 0; i < cnt; ++i) distances_host[(int64_t)node * E + i] = dst[i];
        }
    };
    int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min(nt, M));
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(worker);
    worker();
    for (auto &t : pool) t.join();
    OFX_REQUIRE(bad_vertex.load() < 0, "ofx_compute_edges_geodesic_host: reached vertex %d, which is not a valid vertex",
                bad_vertex.load());
    return 0;
}

extern "C" int ofx_node_and_edge_clean_up_host(const int32_t *edges_host, int32_t M, int32_t E, uint8_t *valid_mask_host)
{
    OFX_REQUIRE(M >= 0 && E >= 0, "ofx_node_and_edge_clean_up_host: bad sizes");
    if (M == 0) return 0;
    OFX_REQUIRE(edges_host && valid_mask_host, "ofx_node_and_edge_clean_up_host: null pointer");
    std::vector<uint8_t> removed(M, 0);          // removed by THIS call (nodes that arrive invalid still count as neighbours, :412)
    for (;;) {
        int newly = 0;
        for (int n = 0; n < M; ++n) {
            if (!valid_mask_host[n]) continue;
            int cnt = 0;
            for (int i = 0; i < E; ++i) {
                const int nb = edges_host[(int64_t)n * E + i];
                if (nb == -1) break;                                                          // :408
                if (nb >= 0 && nb < M && removed[nb]) continue;
                ++cnt;
            }
            if (cnt <= 1) { valid_mask_host[n] = 0; removed[n] = 1; ++newly; }                // :419-424
        }
        if (!newly) break;
    }
    return 0;
}

extern "C" int ofx_compute_clusters_host(const int32_t *edges_host, int32_t M, int32_t E, int32_t *clusters_host,
                                         int32_t *sizes_host, int32_t *n_clusters_host)
{
    OFX_REQUIRE(M >= 0 && E >= 0 && n_clusters_host, "ofx_compute_clusters_host: bad arguments");
    *n_clusters_host = 0;
    if (M == 0) return 0;
    OFX_REQUIRE(edges_host && clusters_host, "ofx_compute_clusters_host: null pointer");
    // union-find over the undirected edges; clusters are numbered by their smallest node (what the reference's traversal
    // in node order yields, :465-472)
    std::vector<int> parent(M);
    for (int n = 0; n < M; ++n) parent[n] = n;
    auto find = [&](int x) { while (parent[x] != x) { parent[x] = parent[parent[x]]; x = parent[x]; } return x; };
    for (int n = 0; n < M; ++n)
        for (int i = 0; i < E; ++i) {
            const int nb = edges_host[(int64_t)n * E + i];
            if (nb == -1) break;                                                              // :455
            OFX_REQUIRE(nb >= 0 && nb < M, "ofx_compute_clusters_host: edge %d -> %d out of range", n, nb);
            const int a = find(n), b = find(nb);
            if (a != b) parent[std::max(a, b)] = std::min(a, b);
        }
    std::vector<int> label(M, -1);
    int nc = 0;
    for (int n = 0; n < M; ++n) {
        const int r = find(n);                   // r <= n: the smallest node of the component
        if (label[r] < 0) { label[r] = nc; if (sizes_host) sizes_host[nc] = 0; ++nc; }
        clusters_host[n] = label[r];
        if (sizes_host) sizes_host[label[r]] += 1;
    }
    *n_clusters_host = nc;
    return 0;
}

// One level of the node pyramid (fusion_with_occlusion/embedded_deformation_graph.py:279-298): nodes in index order; a node
// joins the coarser level unless a node kept before it is closer than `coverage`; up_idx[i] = position (within the kept
// list) of the kept node nearest to node i at the moment i is examined (0 for the first node).  float32 norms
// sqrtf((dx^2+dy^2)+dz^2), first index among equal norms, float64 comparison with the bound.
extern "C" int ofx_downsample_nodes_host(const float *nodes_host, int32_t M, double coverage, int32_t *down_idx_host,
                                         int32_t *n_down_host, int32_t *up_idx_host)
{
    OFX_REQUIRE(M >= 0 && n_down_host, "ofx_downsample_nodes_host: bad arguments");
    *n_down_host = 0;
    if (M == 0) return 0;
    OFX_REQUIRE(nodes_host && down_idx_host && up_idx_host, "ofx_downsample_nodes_host: null pointer");
    int kept = 0;
    for (int i = 0; i < M; ++i) {
        if (kept == 0) { up_idx_host[i] = i; down_idx_host[kept++] = i; continue; }          // :280-283
        const float *p = nodes_host + 3 * (int64_t)i;
        float best = INFINITY;
        int bi = 0;
        for (int k = 0; k < kept; ++k) {
            const float *q = nodes_host + 3 * (int64_t)down_idx_host[k];
            const float dx = q[0] - p[0], dy = q[1] - p[1], dz = q[2] - p[2];
            const float d = std::sqrt(((dx * dx) + (dy * dy)) + (dz * dz));
            if (d < best) { best = d; bi = k; }
        }
        up_idx_host[i] = bi;                                                                // :290
        if ((double)best < coverage) continue;                                              // :292
        down_idx_host[kept++] = i;
    }
    *n_down_host = kept;
    return 0;
}
