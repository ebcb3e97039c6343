/* ofx_oracle_graph.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C) of the reference functions on the "next" rows of SURVEY.md section 8
 * (f2 node-insertion machinery, f3 truncated region), each citing the reference file:line it follows
 * (paths relative to /root/reference/):
 *     nearest node                fusion_with_occlusion/warpfield.py:463  (pykdtree query, k = 1, float32)
 *     greedy node insertion       fusion_with_occlusion/embedded_deformation_graph.py:516-536
 *     sample_nodes                csrc/cpu/graph_proc.cpp:79-136  (randomShuffle = false, as edg.py:196 calls it)
 *     erode_mesh                  csrc/cpu/graph_proc.cpp:17-77
 *     compute_truncated_region    fusion_with_occlusion/tsdf.py:706-745
 * Only tests/ and __graft_entry__.smoke() may load this; the product never does.
 *
 * Parity pin (tests/test_oracle_golden_graph.py, fixtures from oracle/make_golden_graph.py):
 *   - greedy insertion against the reference's own EDGraph.update (Python, imported in place);
 *   - sample_nodes / erode_mesh against the reference's own C++ compiled from /root/reference into
 *     oracle/_ref/ref_graph.so (oracle/build_ref.sh), including adversarial vertex pairs whose squared
 *     distance straddles coverage^2 depending on the summation order (pins Eigen's a0 + (a1 + a2));
 *   - truncated region against the reference's own numba function.
 * Build: -ffp-contract=off (no implicit fused multiply-adds).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* numpy / pykdtree float32 order: ((dx*dx) + (dy*dy)) + (dz*dz)   (np.linalg.norm axis=1 = sqrt(add.reduce(x*x))) */
static inline float d2_seq(const float *a, const float *b)
{
    const float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return ((dx * dx) + (dy * dy)) + (dz * dz);
}

/* Eigen 3 Vector3f::squaredNorm(): not vectorised (3 floats), completely unrolled redux that splits the
 * range in halves: [0,1) + [1,3)  ->  dx*dx + (dy*dy + dz*dz)   (Eigen/src/Core/Redux.h, redux_novec_unroller) */
static inline float d2_eigen(const float *a, const float *b)
{
    const float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return (dx * dx) + ((dy * dy) + (dz * dz));
}

/* warpfield.py:463  dist, nearest_node = kdtree.query(points, k=1): float32 distances, ties -> lower index */
void ofx_oracle_nearest_node(const float *pts, int64_t n, const float *nodes, int M, float *dist, int32_t *idx)
{
    for (int64_t i = 0; i < n; ++i) {
        float best = INFINITY;
        int bi = -1;
        for (int j = 0; j < M; ++j) {
            const float d2 = d2_seq(pts + 3 * i, nodes + 3 * j);
            if (d2 < best) { best = d2; bi = j; }
        }
        dist[i] = sqrtf(best);
        idx[i] = bi;
    }
}

/* embedded_deformation_graph.py:516-536.  Candidates in the given order; a candidate is dropped when some
 * existing node (:522) or some already accepted vertex (:530-531) is closer than node_coverage:
 * np.min(np.linalg.norm(..., axis=1)) < node_coverage with float32 norms and a Python-float bound, i.e.
 * (double)sqrtf(d2) < coverage.  Returns the number of accepted vertices, ids in accepted[]. */
int ofx_oracle_greedy_update(const float *verts, const int32_t *cand, int n_cand, const float *nodes, int M,
                             double coverage, int32_t *accepted)
{
    int na = 0;
    for (int c = 0; c < n_cand; ++c) {
        const float *v = verts + 3 * (int64_t)cand[c];
        int covered = 0;
        for (int j = 0; j < M && !covered; ++j) covered = (double)sqrtf(d2_seq(nodes + 3 * j, v)) < coverage;
        for (int a = 0; a < na && !covered; ++a)
            covered = (double)sqrtf(d2_seq(verts + 3 * (int64_t)accepted[a], v)) < coverage;
        if (!covered) accepted[na++] = cand[c];
    }
    return na;
}

/* graph_proc.cpp:79-136 with randomShuffle = false: vertices in index order; skipped when eroded (:115);
 * rejected when squaredNorm(point - node) <= nodeCoverage^2 (float32, :121) for an accepted node. */
int ofx_oracle_sample_nodes(const float *verts, const uint8_t *non_eroded, int V, float coverage,
                            int use_only_non_eroded, int32_t *node_idx, float *node_pos)
{
    const float c2 = coverage * coverage;                               /* :89 */
    int na = 0;
    for (int v = 0; v < V; ++v) {
        if (use_only_non_eroded && !non_eroded[v]) continue;
        int is_node = 1;
        for (int a = 0; a < na && is_node; ++a)
            if (d2_eigen(verts + 3 * (int64_t)v, verts + 3 * (int64_t)node_idx[a]) <= c2) is_node = 0;
        if (is_node) {
            node_idx[na] = v;
            if (node_pos) memcpy(node_pos + 3 * na, verts + 3 * (int64_t)v, 12);
            ++na;
        }
    }
    return na;
}

/* graph_proc.cpp:17-77: n_iter rounds of "count the faces at every vertex, keep the faces whose three
 * vertices all have >= min_neighbours faces"; vertices of surviving faces are non-eroded. */
void ofx_oracle_erode_mesh(const int32_t *faces, int F, int V, int n_iter, int min_neighbours, uint8_t *non_eroded)
{
    uint8_t *alive = (uint8_t *)malloc(F > 0 ? F : 1);
    int *cnt = (int *)malloc(sizeof(int) * (V > 0 ? V : 1));
    memset(alive, 1, F);
    for (int it = 0; it < n_iter; ++it) {
        memset(cnt, 0, sizeof(int) * V);
        for (int f = 0; f < F; ++f)
            if (alive[f]) { cnt[faces[3 * f]]++; cnt[faces[3 * f + 1]]++; cnt[faces[3 * f + 2]]++; }
        for (int f = 0; f < F; ++f)
            if (alive[f] && !(cnt[faces[3 * f]] >= min_neighbours && cnt[faces[3 * f + 1]] >= min_neighbours &&
                              cnt[faces[3 * f + 2]] >= min_neighbours))
                alive[f] = 0;
    }
    memset(non_eroded, 0, V);
    for (int f = 0; f < F; ++f)
        if (alive[f]) { non_eroded[faces[3 * f]] = 1; non_eroded[faces[3 * f + 1]] = 1; non_eroded[faces[3 * f + 2]] = 1; }
    free(alive);
    free(cnt);
}

/* tsdf.py:706-745.  A voxel is in the truncated region unless |tsdf| > 0.9 (:724), it lies on the volume
 * boundary (:729), or one of its 26 neighbours differs by more than max_diff (:740: float32 difference,
 * float64 comparison).  The write at :737-738 is dead code (its condition is the one :724 already
 * `continue`d on), so the prange race SURVEY.md notes cannot fire and the result is deterministic.
 * 0.9 is a float64 literal: numba compares float64(abs(float32)) > 0.9. */
void ofx_oracle_truncated_region(const float *tsdf, int W, int H, int D, double max_diff, uint8_t *out)
{
#pragma omp parallel for schedule(static)
    for (int w = 0; w < W; ++w)
        for (int h = 0; h < H; ++h)
            for (int d = 0; d < D; ++d) {
                const int64_t i = ((int64_t)w * H + h) * D + d;
                const float t = tsdf[i];
                uint8_t r = 1;
                if ((double)fabsf(t) > 0.9) r = 0;
                else if (d == 0 || d == D - 1 || h == 0 || h == H - 1 || w == 0 || w == W - 1) r = 0;
                else
                    for (int dw = -1; dw < 2 && r; ++dw)
                        for (int dh = -1; dh < 2 && r; ++dh)
                            for (int dd = -1; dd < 2; ++dd) {
                                const float nb = tsdf[((int64_t)(w + dw) * H + (h + dh)) * D + (d + dd)];
                                if ((double)fabsf(nb - t) > max_diff) { r = 0; break; }
                            }
                out[i] = r;
            }
}
