/* ofx.h -- C ABI of libofx.so: the B200 (sm_100a) replacement for the native entry points of
 * OcclusionFusion's non-rigid fusion hot path.
 *
 * What it replaces in the reference (paths relative to the reference root):
 *   - the PyCUDA `SourceModule(...).get_function("integrate")` launch
 *         fusion_with_occlusion/tsdf.py:192-288 (kernel), :290 (binding), :405-441 (launch)
 *   - the numba.cuda launch `deform_lbs_cuda[grid, block](...)`
 *         fusion_with_occlusion/warpfield.py:234-266, 288-294, 607-630
 *   - the torch warp `Registration.deform_ED`      NonRigidICP/model/registration_fusion.py:157-184
 *   - the pykdtree query + numpy weights in `WarpField.skin`   fusion_with_occlusion/warpfield.py:83-129
 *   - the numpy/numba CPU branch of `TSDFVolume.integrate`     fusion_with_occlusion/tsdf.py:442-494
 *     (this branch defines the arithmetic: see DESIGN.md "contract arithmetic")
 *   - the callers either side of that path (declared further down, each with its own citation):
 *         graph maintenance      NeuralNRT._C (pybind11, csrc/main.cpp:7-40): erode_mesh, sample_nodes, compute_edges_geodesic,
 *                                node_and_edge_clean_up, compute_clusters, compute_pixel_anchors_euclidean
 *                                (csrc/cpu/graph_proc.cpp), the Python loops of EDGraph.update and find_unreachable_nodes
 *         surface extraction     numba compute_truncated_region (tsdf.py:706-745), skimage marching cubes (tsdf.py:794)
 *
 * Conventions
 *   - plain C: opaque handle, raw pointers, sizes; no C++ or torch types; no exceptions.
 *   - every function returns 0 on success, non-zero on failure; ofx_last_error() returns the
 *     message of the calling thread's last failure.
 *   - pointers are DEVICE pointers owned by the caller (torch tensors on the Python side)
 *     unless the parameter name ends in `_host`.
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing
 *     synchronises except where stated.
 *   - one handle per GPU (per volume shard); the handle's device must be current.
 *   - there is no CPU fallback: without a CUDA device every call that would launch fails.
 *
 * Volume layout (tsdf.py:133-141, 294-304): three float32 arrays of shape (nx_local, Dy, Dz),
 * C order, linear index (lx*Dy + iy)*Dz + iz, iz fastest.  A shard owns the x-planes
 *     gx = (lx / x_block) * x_block * n_shards + shard * x_block + lx % x_block
 * (x_block = Dx / n_shards gives contiguous slabs; smaller blocks deal planes round-robin).
 */
#ifndef OFX_H
#define OFX_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OFX_ABI_VERSION 1
#define OFX_MAX_K 8

typedef struct ofx_volume ofx_volume; /* opaque */

typedef struct ofx_volume_desc {
    int32_t dims[3];      /* global Dx, Dy, Dz                                   tsdf.py:95-118 */
    int32_t n_shards;     /* 1 = whole volume on this GPU                                        */
    int32_t shard;        /* this GPU's shard index in [0, n_shards)                             */
    int32_t x_block;      /* planes per dealt block; Dx % (x_block*n_shards) == 0               */
    float origin[3];      /* _vol_origin (float32)                               tsdf.py:129    */
    double voxel_size;    /* _voxel_size (float64 in the reference)              tsdf.py:106    */
    double trunc_margin;  /* _trunc_margin                                       tsdf.py:127    */
    float *tsdf;          /* (nx_local, Dy, Dz) device arrays, caller-owned       tsdf.py:133-141 */
    float *weight;
    float *colour;
} ofx_volume_desc;

/* counters written by ofx_integrate (device array of OFX_N_STATS uint64) */
enum {
    OFX_STAT_UPDATED = 0,    /* voxels whose tsdf/weight/colour were written                  */
    OFX_STAT_SKIN_VALID = 1, /* voxels with K anchors inside 4*coverage that were visited     */
    OFX_STAT_BRICKS = 2,     /* bricks visited                                                */
    OFX_STAT_WEIGHT_SUM = 3, /* sum of the new weights of updated voxels (integer-valued)     */
    OFX_N_STATS = 4
};

int ofx_abi_version(void);
const char *ofx_last_error(void);

/* ---- volume handle ------------------------------------------------------------------------ */
int ofx_volume_create(ofx_volume **out, const ofx_volume_desc *desc);
int ofx_volume_destroy(ofx_volume *v);
/* fills tsdf=1, weight=0, colour=0                                         tsdf.py:133-141 */
int ofx_volume_reset(ofx_volume *v, void *stream);
int64_t ofx_volume_num_voxels(const ofx_volume *v); /* local voxels: nx_local*Dy*Dz */

/* Graph nodes (canonical positions g, EDGraph.nodes) -- call again after node insertion
 * (WarpField.update_graph, warpfield.py:487-582 -> :532-533 rebuild tree + re-skin).
 * Rebuilds the per-voxel skin cache (anchor ids + normalised weights; the device form of WarpField.skin_tsdf's cache,
 * warpfield.py:131-141).  K = graph_neighbours (warpfield.py:60), 1..OFX_MAX_K, K <= M.
 * Waits on `stream` once in the middle (the sizes of the cache it has to provide); the results of its last kernels -- the
 * number of active bricks, the brick-node overflow error, the statistics the per-frame launch configuration is chosen
 * from -- are read by the next call on the handle that needs them, which reports the overflow if there was one. */
int ofx_set_graph(ofx_volume *v, const float *nodes, int32_t M, double node_coverage, int32_t K,
                  void *stream);
/* Node transforms in origin form: rotations (M,3,3) row-major, translations T = -R g + g + t
 * (WarpField.update_transformations, warpfield.py:389-420). */
int ofx_set_transforms(ofx_volume *v, const float *rotations, const float *translations,
                       void *stream);

/* One frame of TSDFVolume.integrate (tsdf.py:378-494): warp every voxel (unless
 * identity_warp != 0: the source frame, tsdf.py:395-398), project, test, update in place.
 * depth / colour_packed are the (H,W) images TSDFVolume.update produces (tsdf.py:545-572);
 * intr_host = {fx, fy, cx, cy} already cast to float32 (tsdf.py:357); camera pose is I (:392).
 * stats (may be NULL) receives OFX_N_STATS counters, overwritten. */
int ofx_integrate(ofx_volume *v, const float *depth, const float *colour_packed, int32_t H,
                  int32_t W, const float *intr_host, double obs_weight, int32_t identity_warp,
                  uint64_t *stats, void *stream);

/* ofx_set_transforms + ofx_integrate(identity_warp = 0) as ONE call: WarpField.update_transformations followed by
 * TSDFVolume.integrate of the registered frame (fusion_with_occlusion/fusion.py:176-179).  The node transforms are
 * written by the frame-preparation launch, so a warped frame costs three kernel launches instead of four; results
 * are identical to the two separate calls. */
int ofx_integrate_warped(ofx_volume *v, const float *rotations, const float *translations,
                         const float *depth, const float *colour_packed, int32_t H, int32_t W,
                         const float *intr_host, double obs_weight, uint64_t *stats, void *stream);

/* Optional look-ahead for frame streams (new in this library; the reference integrates strictly frame by frame):
 * stages the NEXT warped frame's node transforms (as ofx_set_transforms) and its list of bricks inside the
 * camera frustum in a second per-frame set, so that it may run on another stream while the current frame
 * integrates.  The next non-identity ofx_integrate (same H, W, intr_host) switches to that set and launches only
 * the fused kernel.  It writes the set that the most recently ISSUED ofx_integrate does not use, so the host calls
 * go integrate(t), prepare(t+1), integrate(t+1), ...; on the device the caller orders prepare(t+1) after
 * integrate(t-1) has finished and integrate(t+1) after prepare(t+1) (events).  Results are identical to
 * ofx_set_transforms + ofx_integrate. */
int ofx_prepare_frame(ofx_volume *v, const float *rotations, const float *translations, int32_t H,
                      int32_t W, const float *intr_host, void *stream);
/* No reference counterpart (the reference integrates one frame at a time on one device).  EVERYTHING of the next warped
 * frame except the fused kernel -- interleaved frame image, depth tiles, node transforms, brick cull with the depth-range
 * test -- into the frame set that is not in use, on `stream`, while the current frame integrates on another stream; the
 * caller orders the call after the frame BEFORE the current one (whose set it overwrites).  The next ofx_integrate must be
 * given the same depth / colour_packed pointers and camera; it launches the fused kernel only (on the full persistent grid:
 * leaving a CTA slot per SM free for the staging launches was measured and costs more than it hides). */
int ofx_stage_frame(ofx_volume *v, const float *rotations, const float *translations, const float *depth,
                    const float *colour_packed, int32_t H, int32_t W, const float *intr_host, void *stream);

/* Materialise what the reference keeps on the host, on request only (vis / tests):
 *   world_pts (n,3)                                   tsdf.py:307
 *   world_anchors (n,K) i32, world_weights (n,K) f32, world_valid_pts (n,) u8   warpfield.py:131-141
 *   deform_tsdf(): warped (n,3) f32 + valid (n,) u8   warpfield.py:369-380
 * Any output pointer may be NULL. n = ofx_volume_num_voxels. */
int ofx_volume_world_points(ofx_volume *v, float *out_pts, void *stream);
int ofx_volume_skin(ofx_volume *v, int32_t *anchors, float *weights, uint8_t *valid, void *stream);
int ofx_volume_deform(ofx_volume *v, float *out_pts, uint8_t *valid, void *stream);

/* ---- free-standing point-set operators ------------------------------------------------------ */
/* TSDFVolume.update (tsdf.py:545-572): im (C,H,W) f32, C >= 4 -> packed colour (H,W) from
 * channels 0..2 (RGB in [0,1]) and depth (H,W) = last channel (im[-1], tsdf.py:554).  The
 * reference frame has C = 6 (RGB + XYZ); the host may upload only RGB + Z (C = 4). */
int ofx_pack_frame(const float *im, int32_t C, int32_t H, int32_t W, float *colour_packed,
                   float *depth, void *stream);
/* WarpField.skin (warpfield.py:83-129): anchors (n,K) i32 (-1 beyond 4*coverage), weights (n,K)
 * f32, valid (n,) u8. */
int ofx_skin(const float *points, int64_t n, const float *nodes, int32_t M, double node_coverage,
             int32_t K, int32_t *anchors, float *weights, uint8_t *valid, void *stream);
/* WarpField.deform (warpfield.py:270-305): out = sum_k w_k (R_k p + T_k) for valid points,
 * p itself otherwise. */
int ofx_warp_points(const float *points, int64_t n, const int32_t *anchors, const float *weights,
                    const uint8_t *valid, int32_t K, const float *rotations,
                    const float *translations, float *out, void *stream);
/* TSDFVolume.check_visibility (tsdf.py:599-612): valid (n,) u8, depth_diff (n,) f64. */
int ofx_check_visibility(const float *points, int64_t n, const float *depth, int32_t H, int32_t W,
                         const float *intr_host, double trunc_margin, uint8_t *valid,
                         double *depth_diff, void *stream);

/* ---- graph maintenance and surface mask (callers either side of the path when nodes are inserted) -------- */
/* source_frame_kdtree.query(points, k=1) (fusion_with_occlusion/warpfield.py:463, WarpField.find_unreachable_nodes):
 * float32 distance to and index of the nearest node, ties -> lower index.  dist or idx may be NULL. */
int ofx_nearest_node(const float *points, int64_t n, const float *nodes, int32_t M, float *dist,
                     int32_t *idx, void *stream);
/* compute_pixel_anchors_euclidean (csrc/cpu/graph_proc.cpp:610-709, binding csrc/main.cpp:35): for every pixel of the
 * (3,H,W) point image with z > 0 its 4 (GRAPH_K, graph_proc.h:8) nearest nodes by squared distance -- among equal distances
 * the later node first -- and their Gaussian weights normalised by their sum; anchors (H,W,4) i32 (-1 = none), weights
 * (H,W,4) f32.  The per-voxel skinning of the fusion path follows WarpField.skin instead (ofx_skin). */
int ofx_pixel_anchors_euclidean(const float *nodes, int32_t M, const float *point_image, int32_t H,
                                int32_t W, float node_coverage, int32_t *anchors, float *weights,
                                void *stream);
/* np.argmin(np.linalg.norm(points - q, axis=1)) for every query q (EDGraph.update re-anchoring the nodes on the new mesh,
 * fusion_with_occlusion/embedded_deformation_graph.py:545-547, which builds the dense (queries x points) matrix on the
 * host): float32 norms, first index among equal norms.  dist or idx may be NULL. */
int ofx_argmin_norm(const float *queries, int64_t n, const float *points, int32_t M, float *dist,
                    int32_t *idx, void *stream);
/* The greedy coverage rule of EDGraph.update (fusion_with_occlusion/embedded_deformation_graph.py:516-536):
 * walk cand[0..n_cand) (vertex ids into verts (V,3), in the caller's order -- farthest first, warpfield.py:469-483);
 * accept a vertex unless an existing node (nodes (M,3), M may be 0) or an already accepted vertex is closer than
 * node_coverage (float32 norm, float64 comparison).  accepted (capacity n_cand) receives the accepted vertex ids in
 * acceptance order, *n_accepted (device) their number.  Identical to the serial loop, order included. */
int ofx_greedy_insert(const float *verts, int32_t V, const int32_t *cand, int32_t n_cand,
                      const float *nodes, int32_t M, double node_coverage, int32_t *accepted,
                      int32_t *n_accepted, void *stream);
/* sample_nodes (csrc/cpu/graph_proc.cpp:79-136, binding csrc/main.cpp:29) with randomShuffle = false, as
 * EDGraph.create_graph_from_mesh calls it (embedded_deformation_graph.py:196-206): vertices in index order, eroded
 * ones skipped when use_only_non_eroded, a vertex becomes a node unless squaredNorm(v - node) <= coverage^2 (float32)
 * for an earlier node.  node_idx (capacity V), node_pos (capacity V x 3, may be NULL), *n_nodes (device). */
int ofx_sample_nodes(const float *verts, int32_t V, const uint8_t *non_eroded, float node_coverage,
                     int32_t use_only_non_eroded, int32_t *node_idx, float *node_pos,
                     int32_t *n_nodes, void *stream);
/* erode_mesh (csrc/cpu/graph_proc.cpp:17-77, binding csrc/main.cpp:28): n_iterations rounds of "keep the faces whose
 * three vertices each touch >= min_neighbours kept faces"; non_eroded (V,) u8 = vertex of a surviving face.
 * Faces (F,3) i32; a face with an index outside [0,V) is ignored (undefined behaviour in the reference). */
int ofx_erode_mesh(const int32_t *faces, int32_t F, int32_t V, int32_t n_iterations,
                   int32_t min_neighbours, uint8_t *non_eroded, void *stream);
/* TSDFVolume.compute_truncated_region (fusion_with_occlusion/tsdf.py:706-745): region (Dx,Dy,Dz) u8, the mask the
 * reference hands to marching cubes (tsdf.py:794).  tsdf is a whole (unsharded) volume, C order. */
int ofx_truncated_region(const float *tsdf, int32_t Dx, int32_t Dy, int32_t Dz, double max_diff,
                         uint8_t *region, void *stream);

/* ---- marching cubes (the scikit-image call at fusion_with_occlusion/tsdf.py:794 and the colour unpack at :797-807) ----
 * Classic form with the case table of csrc/ofx_mc_tables.h, ONE call, no host read inside.  vol / mask / colour are whole
 * (Dx,Dy,Dz) volumes, C order (mask, colour may be NULL); cube i = (x*Dy + y)*Dz + z is processed when mask == NULL or
 * mask[i] != 0; corners with value < level are inside.  Outputs (DEVICE):
 *   counts (2,) i32   [0] vertices, [1] triangles of the mesh -- always written;
 *   verts (cap_verts,3) f32   one vertex per crossed grid edge of a processed cube, ascending by edge
 *                     (3 * linear index of the edge's lower corner + axis), in voxel coordinates: lower corner +
 *                     (level - v0)/(v1 - v0) along the axis, float32;
 *   normals (cap_verts,3) f32 or NULL   unit gradient of vol at the vertex (towards increasing values);
 *   rgb (cap_verts,3) u8 or NULL        colour of the nearest voxel unpacked to r, g, b (tsdf.py:797-807; needs colour);
 *   faces (cap_tris,3) i32    vertex indices, triangles in cube order.
 * When counts[0] > cap_verts or counts[1] > cap_tris nothing but counts is written: read counts, come back with room
 * (cap 0 / NULL buffers = count only).  Dz <= 21845. */
int ofx_mc_mesh(const float *vol, const uint8_t *mask, const float *colour, int32_t Dx, int32_t Dy, int32_t Dz,
                float level, int32_t cap_verts, int32_t cap_tris, float *verts, float *normals, uint8_t *rgb,
                int32_t *faces, int32_t *counts, void *stream);

/* ---- the edge half of the graph builder on the device ---------------------------------------------------------
 * compute_edges_geodesic (csrc/cpu/graph_proc.cpp:155-300, binding csrc/main.cpp:30; callers
 * embedded_deformation_graph.py:226-233, 313-320, 575-582): for every node (node_idx: its mesh vertex, < 0 = skip) the
 * first max_neighbours OTHER nodes in order of geodesic distance over the mesh.  DEVICE pointers: verts (V,3) f32, valid
 * (V,) u8 (may be NULL unless only_valid), faces (F,3) i32, node_idx (M,) i32; outputs edges (M, max_neighbours) i32 and
 * distances (same shape, f32) are filled for the neighbours found (pre-fill -1 / 0 as the reference's callers do),
 * n_found (M,) i32 their number.  enforce_total == 0 limits the walk to 2 * node_coverage.  Identical to the reference's
 * search, ties included (nodes whose answer depends on the order of equal distances are replayed with the reference's
 * heap, step by step).  Synchronises the stream.  status_host (2 ints, may be NULL): [0] 0 = ok, [1] nodes replayed.
 * The edge WEIGHTS follow from the distances on the host: ofx_edge_weights_host. */
int ofx_compute_edges_geodesic(const float *verts, const uint8_t *valid, int32_t V, const int32_t *faces, int32_t F,
                               const int32_t *node_idx, int32_t M, int32_t max_neighbours, float node_coverage,
                               int32_t *edges, float *distances, int32_t *n_found, int32_t only_valid,
                               int32_t enforce_total, int32_t *status_host, void *stream);
/* The same for several searches over ONE mesh in one launch (the levels of EDGraph.create_graph_pyramid,
 * embedded_deformation_graph.py:300-320: their node sets, neighbour counts and radii differ, the mesh does not): job j owns
 * job_count_host[j] consecutive entries of node_idx and the next job_count_host[j] * job_neighbours_host[j] entries of edges /
 * distances; job_flags_host[j]: bit 0 = only_valid, bit 1 = enforce_total.  At most 8 jobs. */
int ofx_compute_edges_geodesic_batch(const float *verts, const uint8_t *valid, int32_t V, const int32_t *faces, int32_t F,
                                     const int32_t *node_idx, int32_t n_jobs, const int32_t *job_count_host,
                                     const int32_t *job_neighbours_host, const float *job_coverage_host,
                                     const int32_t *job_flags_host, int32_t *edges, float *distances, int32_t *n_found,
                                     int32_t *status_host, void *stream);
/* exp(-d^2 / (2 c^2)) normalised per node (graph_proc.cpp:150-152, 281-296) from the distances above: HOST arrays (the
 * reference's weights go through the C library's expf; so do these). */
int ofx_edge_weights_host(const float *distances_host, const int32_t *n_found_host, int32_t M, int32_t E,
                          float node_coverage, float *weights_host);
/* node_and_edge_clean_up (graph_proc.cpp:388-438, binding main.cpp:32) on DEVICE arrays: clears valid_mask (M,) u8 of every
 * node left with at most one neighbour, repeatedly (the fixed point does not depend on the sweep order). */
int ofx_node_and_edge_clean_up(const int32_t *edges, int32_t M, int32_t E, uint8_t *valid_mask, void *stream);
/* compute_clusters (graph_proc.cpp:440-481, binding main.cpp:33) on DEVICE arrays: connected components of the undirected
 * edge set, numbered by their smallest node; clusters (M,) i32, sizes (M,) i32 (first *n_clusters entries), *n_clusters
 * (device).  Synchronises the stream (error check). */
int ofx_compute_clusters(const int32_t *edges, int32_t M, int32_t E, int32_t *clusters, int32_t *sizes,
                         int32_t *n_clusters, void *stream);
/* ofx_compute_clusters without any allocation.  NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The label scratch of the kernel lives in a file-scope device array of OFX_NOALLOC_MAX_NODES ints (256 KiB, held by every
 * GPU that loads the library until exit); more nodes are refused with an error (use ofx_compute_clusters).  Same outputs as
 * ofx_compute_clusters; the range check of the edges is reported through *status (DEVICE int32: 0 = ok, n + 1 = node n
 * has an edge out of range) instead of a host read, so the call neither allocates nor synchronises.  All calls on one GPU
 * share the array: calls that overlap on the GPU (different streams) race -- order them. */
#define OFX_NOALLOC_MAX_NODES 65536
int ofx_compute_clusters_noalloc(const int32_t *edges, int32_t M, int32_t E, int32_t *clusters, int32_t *sizes,
                                 int32_t *n_clusters, int32_t *status, void *stream);
/* up_sample_idx of one pyramid level (embedded_deformation_graph.py:279-298) on DEVICE arrays: for every node the position,
 * within down_idx (the kept nodes, ascending: ofx_greedy_insert over the nodes in index order), of the nearest node kept
 * before it. */
int ofx_upsample_index(const float *nodes, int32_t M, const int32_t *down_idx, int32_t n_down, int32_t *up_idx,
                       void *stream);

/* ---- host half of the graph builder (HOST pointers; no device work) ------------------------------------------ */
/* compute_edges_geodesic (csrc/cpu/graph_proc.cpp:155-300, binding csrc/main.cpp:30; callers
 * embedded_deformation_graph.py:226-233, 313-320): for every node (node_idx: its mesh vertex, < 0 = skip) walk the mesh
 * outwards in order of geodesic distance until max_neighbours other nodes were met; edges (M, max_neighbours) i32 and
 * weights / distances (same shape, f32) are filled for the neighbours found (the caller pre-fills -1 / 0 / 0 as the
 * reference does), node_to_vertex (M, V) f32 receives the distance of every vertex visited (may be NULL).
 * only_valid: vertices with valid[v] == 0 are never entered (reaching one is an error; the reference exits the process).
 * enforce_total == 0 limits the walk to 2 * node_coverage.  n_threads <= 0: all host threads (nodes are independent). */
int ofx_compute_edges_geodesic_host(const float *verts_host, const uint8_t *valid_host, int32_t V,
                                    const int32_t *faces_host, int32_t F, const int32_t *node_idx_host,
                                    int32_t M, int32_t max_neighbours, float node_coverage,
                                    int32_t *edges_host, float *weights_host, float *distances_host,
                                    float *node_to_vertex_host, int32_t only_valid,
                                    int32_t enforce_total, int32_t n_threads);
/* node_and_edge_clean_up (graph_proc.cpp:388-438, binding main.cpp:32): clears valid_mask (M,) u8 of every node left
 * with at most one neighbour, repeatedly, in node order. */
int ofx_node_and_edge_clean_up_host(const int32_t *edges_host, int32_t M, int32_t E,
                                    uint8_t *valid_mask_host);
/* compute_clusters (graph_proc.cpp:440-481, binding main.cpp:33): connected components of the undirected edge set,
 * numbered by their smallest node; clusters (M,) i32, sizes (capacity M, may be NULL), *n_clusters. */
int ofx_compute_clusters_host(const int32_t *edges_host, int32_t M, int32_t E, int32_t *clusters_host,
                              int32_t *sizes_host, int32_t *n_clusters_host);

/* One level of EDGraph.create_graph_pyramid's greedy down-sampling (embedded_deformation_graph.py:279-298): down_idx
 * (capacity M) = nodes kept for the coarser level, in order; up_idx (M,) = for every node the position, within the kept
 * list, of the nearest node kept before it. */
int ofx_downsample_nodes_host(const float *nodes_host, int32_t M, double coverage, int32_t *down_idx_host,
                              int32_t *n_down_host, int32_t *up_idx_host);

/* ---- multi-GPU plumbing ------------------------------------------------------------------------ */
/* Copies `bytes` (a multiple of 16, both pointers 16-byte aligned) from `src` to `dst` with `ctas`
 * thread blocks.  `src` may be a PEER pointer (another GPU's buffer mapped into this process through
 * NVLink / NVSwitch, e.g. torch symmetric memory): every rank pulls the per-frame packet
 * [R | T | colour | depth] from the rank that received the camera frame.  New in this library (the
 * reference has no multi-device path); replaces the NCCL broadcast on the per-frame path. */
int ofx_peer_copy(void *dst, const void *src, int64_t bytes, int32_t ctas, void *stream);
/* The frame packet pushed ONCE through an NVSwitch multicast mapping (no reference counterpart: the reference is single-device):
 * copies `bytes` (multiple of 16) from `src` (local device memory) to `mc_dst`, a multicast address that the switch
 * replicates into the same offset of every participating GPU's buffer (torch symmetric memory: handle.multicast_ptr), with
 * multimem.st.  Completion on the peers is the caller's hand-shake (a barrier after this launch in stream order). */
int ofx_multicast_copy(void *mc_dst, const void *src, int64_t bytes, int32_t ctas, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* OFX_H */
