"""WarpField -- drop-in for the reference class of the same name
(/root/reference/fusion_with_occlusion/warpfield.py:21-605) on the fusion hot path: skinning
(K nearest nodes + Gaussian weights), linear-blend warp, transform bookkeeping, node insertion.

Host state (`rotations`, `translations`, `deformed_nodes`, `frame_id`, ...) keeps the
reference's names, shapes and update rules; device state (node table, anchor cache) lives
behind the TSDFVolume's libofx handle and is refreshed lazily before the next integrate.
"""
import logging
import os

import numpy as np
import torch

from . import _capi


def _rot_apply(R, g):
    """np.array([R[i] @ g[i] for i in range(N)]) as one batched matmul (bit-identical results)."""
    R, g = np.asarray(R), np.asarray(g)
    return np.matmul(R, g[:, :, None])[:, :, 0]


class WarpField:
    def __init__(self, graph, tsdf, visualizer=None, kdtree_leaf_size=16):
        """As warpfield.py:22-75.  graph: object with `.nodes` (M,3) float32 and
        `.graph_generation_parameters["node_coverage"]`; tsdf: occlusionfusion_b200.TSDFVolume."""
        _capi.require_cuda()
        self._lib = _capi.load()
        self.graph = graph
        self.tsdf = tsdf
        self.vis = visualizer
        self.frame_id = self.tsdf.fopt.source_frame
        self.log = logging.getLogger(__name__)
        self.kdtree_leaf_size = kdtree_leaf_size        # kept for signature parity; KNN is exact brute force/culling

        N = self.graph.nodes.shape[0]
        self.rotations = np.tile(np.eye(3, dtype=np.float32).reshape((1, 3, 3)), (N, 1, 1))   # warpfield.py:52
        self.translations = np.zeros_like(self.graph.nodes)                                  # warpfield.py:53
        self.deformed_nodes = self.graph.nodes.copy()                                        # warpfield.py:54
        self.updating_warpfield = False
        self.graph_neighbours = min(N, 4)                                                    # warpfield.py:60
        self.node_coverage = self.graph.graph_generation_parameters["node_coverage"]
        self.gpu = True

        fopt = self.tsdf.fopt
        if hasattr(fopt, "datadir") and fopt.datadir is not None:                            # warpfield.py:66-73
            self.savepath = os.path.join(fopt.datadir, "results")
            for sub in ("", "deformed_model", "deformed_nodes", "node_motion"):
                os.makedirs(os.path.join(self.savepath, sub), exist_ok=True)
            np.save(os.path.join(self.savepath, "deformed_nodes", f"{fopt.source_frame}.npy"), self.deformed_nodes)
        else:
            self.savepath = None
        self._graph_key = None          # (id of node array, M, K, coverage) last uploaded
        self._transforms_dirty = True
        self._RT_host = None            # pinned staging [R | T] for the per-frame upload
        self._R_dev = self._T_dev = None

    @property
    def device(self):
        return self.tsdf.device

    # ------------------------------------------------------------------ device sync
    def _sync_graph(self):
        """Upload the node table and rebuild the anchor cache if the graph changed since the last upload."""
        nodes = self.graph.nodes
        key = (id(nodes), nodes.shape[0], int(self.graph_neighbours), float(self.node_coverage))
        if key != self._graph_key or self.updating_warpfield:
            with torch.cuda.device(self.device):
                nd = _capi.to_device(nodes, torch.float32, self.device)
                _capi.check(self._lib.ofx_set_graph(self.tsdf._h, _capi.ptr(nd), nodes.shape[0], float(self.node_coverage),
                                                    int(self.graph_neighbours), _capi.stream_ptr(self.device)))
            self._graph_key = key
            self._transforms_dirty = True
            for a in ("world_anchors", "world_weights", "world_valid_pts"):
                if hasattr(self, a):
                    delattr(self, a)

    def _stage_transforms(self):
        """Host R, T -> device [R (9M) | T (3M)] through one pinned staging buffer and one asynchronous upload on the
        current stream.  Returns the two device views; the caller hands them to the library."""
        nodes = self.graph.nodes
        M = nodes.shape[0]
        R = np.asarray(self.rotations)
        T = np.asarray(self.translations)
        assert R.shape[0] == M == T.shape[0], \
            f"Nodes mismatch between transformation data:{self.rotations.shape} and graph nodes:{nodes.shape}"
        if self._RT_host is None or self._RT_host.shape[1] != 12 * M:
            # two staging buffers used in turn: the upload of frame t+1 may be staged while frame t's is in flight
            with torch.cuda.device(self.device):
                self._RT_host = torch.empty((2, 12 * M), dtype=torch.float32, pin_memory=True)
                self._RT_dev = torch.empty((2, 12 * M), dtype=torch.float32, device=self.device)
                self._RT_done = [torch.cuda.Event(), torch.cuda.Event()]
            self._RT_used = [False, False]
            self._RT_i = 0
            # views made once (the frame loop is host-paced on small volumes: every Python object per frame counts)
            self._RT_np = [self._RT_host[k].numpy() for k in range(2)]
            self._RT_views = [(self._RT_dev[k, :9 * M], self._RT_dev[k, 9 * M:]) for k in range(2)]
        i = self._RT_i = 1 - self._RT_i
        if self._RT_used[i]:
            self._RT_done[i].synchronize()       # the upload before last has left this staging buffer
        h = self._RT_np[i]
        h[:9 * M] = R.reshape(-1)                # casts to float32 like the reference's device upload
        h[9 * M:] = T.reshape(-1)
        comp = getattr(self.tsdf, "_comp", None)          # inside TSDFVolume.integrate: the stream it looked up
        if comp is None:
            comp = torch.cuda.current_stream(self.device)
        self._RT_dev[i].copy_(self._RT_host[i], non_blocking=True)      # H2D: runs on the destination device's current stream
        self._RT_done[i].record(comp)
        self._RT_used[i] = True
        self._R_dev, self._T_dev = self._RT_views[i]
        return self._R_dev, self._T_dev

    def _sync_device(self):
        """Upload graph / transforms if they changed since the last integrate."""
        self._sync_graph()
        if self._transforms_dirty:
            R_dev, T_dev = self._stage_transforms()
            with torch.cuda.device(self.device):
                _capi.check(self._lib.ofx_set_transforms(self.tsdf._h, _capi.ptr(R_dev), _capi.ptr(T_dev),
                                                         _capi.stream_ptr(self.device)))
            self._transforms_dirty = False

    # ------------------------------------------------------------------ skinning
    def skin(self, points, nodes=None, ensure_num_neigbours=False):
        """Skinning weights and anchors (warpfield.py:83-129).

        Returns anchors (n,K) int32 (-1 beyond 4*node_coverage), weights (n,K) float32, valid (n,) bool."""
        if nodes is None:
            nodes = self.graph.nodes
        K = int(self.graph_neighbours)
        pts = _capi.to_device(points, torch.float32, self.device).reshape(-1, 3)
        nd = _capi.to_device(nodes, torch.float32, self.device).reshape(-1, 3)
        n = pts.shape[0]
        anchors = torch.empty((n, K), dtype=torch.int32, device=self.device)
        weights = torch.empty((n, K), dtype=torch.float32, device=self.device)
        valid = torch.empty(n, dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _capi.check(self._lib.ofx_skin(_capi.ptr(pts), n, _capi.ptr(nd), nd.shape[0], float(self.node_coverage), K,
                                           _capi.ptr(anchors), _capi.ptr(weights), _capi.ptr(valid),
                                           _capi.stream_ptr(self.device)))
        return anchors.cpu().numpy(), weights.cpu().numpy(), valid.cpu().numpy().astype(bool)

    def skin_tsdf(self):
        """(world_anchors, world_weights, world_valid_pts) over all voxels (warpfield.py:131-141).
        Materialised on request only; integrate() uses the compact device-side anchor cache."""
        self._sync_device()
        if not hasattr(self, "world_anchors") or self.updating_warpfield:
            n, K = int(self._lib.ofx_volume_num_voxels(self.tsdf._h)), int(self.graph_neighbours)
            anchors = torch.empty((n, K), dtype=torch.int32, device=self.device)
            weights = torch.empty((n, K), dtype=torch.float32, device=self.device)
            valid = torch.empty(n, dtype=torch.uint8, device=self.device)
            with torch.cuda.device(self.device):
                _capi.check(self._lib.ofx_volume_skin(self.tsdf._h, _capi.ptr(anchors), _capi.ptr(weights),
                                                      _capi.ptr(valid), _capi.stream_ptr(self.device)))
            self.world_anchors = anchors.cpu().numpy()
            self.world_weights = weights.cpu().numpy()
            self.world_valid_pts = valid.cpu().numpy().astype(bool)
        return self.world_anchors, self.world_weights, self.world_valid_pts

    # ------------------------------------------------------------------ deformation
    def _warp(self, points, anchors, weights, valid_pts, rotations, translations):
        pts = _capi.to_device(points, torch.float32, self.device).reshape(-1, 3)
        n = pts.shape[0]
        a = _capi.to_device(anchors, torch.int32, self.device).reshape(n, -1)
        w = _capi.to_device(weights, torch.float32, self.device).reshape(n, -1)
        v = _capi.to_device(np.asarray(valid_pts).astype(np.uint8) if not isinstance(valid_pts, torch.Tensor)
                            else valid_pts, torch.uint8, self.device).reshape(-1)
        R = _capi.to_device(np.asarray(rotations).astype(np.float32), torch.float32, self.device)
        T = _capi.to_device(np.asarray(translations).astype(np.float32), torch.float32, self.device)
        out = torch.empty_like(pts)
        with torch.cuda.device(self.device):
            _capi.check(self._lib.ofx_warp_points(_capi.ptr(pts), n, _capi.ptr(a), _capi.ptr(w), _capi.ptr(v),
                                                  a.shape[1], _capi.ptr(R), _capi.ptr(T), _capi.ptr(out),
                                                  _capi.stream_ptr(self.device)))
        return out.cpu().numpy()

    def deform(self, points, anchors, weights, reshape_gpu_vol, valid_pts):
        """Linear-blend warp of points (warpfield.py:270-305); invalid points pass through."""
        return self._warp(points, anchors, weights, valid_pts, self.rotations, self.translations)

    def deform_normals(self, normals, anchors, weights, reshape_gpu_vol, valid_pts):
        """Rotate normals by the blended rotations (translation 0) and renormalise (warpfield.py:312-345)."""
        d = self._warp(normals, anchors, weights, valid_pts, self.rotations,
                       np.zeros_like(self.translations, dtype=np.float32))
        return d / np.linalg.norm(d, axis=1, keepdims=True)

    def deform_mesh(self, vertices, normals):
        """Deform canonical mesh vertices (and normals) (warpfield.py:347-367)."""
        skin_anchors, skin_weights, valid_verts = self.skin(vertices)
        reshape_gpu_vol = [vertices.shape[0], 1, 1]
        deformed_vertices = self.deform(vertices, skin_anchors, skin_weights, reshape_gpu_vol, valid_verts)
        deformed_normals = None
        if normals is not None:
            deformed_normals = self.deform_normals(normals, skin_anchors, skin_weights, reshape_gpu_vol, valid_verts)
        return deformed_vertices, deformed_normals, skin_anchors, skin_weights, valid_verts

    def deform_tsdf(self):
        """Deform the TSDF voxel grid (warpfield.py:369-380): (warped (N,3) float32, valid (N,) bool).
        For vis / tests; integrate() never materialises this."""
        assert self.frame_id == self.tsdf.frame_id, \
            f"Warpfield maps to:{self.frame_id}th frame but TSDF maps to:{self.tsdf.frame_id}th frame"
        self._sync_device()
        n = int(self._lib.ofx_volume_num_voxels(self.tsdf._h))
        out = torch.empty((n, 3), dtype=torch.float32, device=self.device)
        valid = torch.empty(n, dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _capi.check(self._lib.ofx_volume_deform(self.tsdf._h, _capi.ptr(out), _capi.ptr(valid),
                                                    _capi.stream_ptr(self.device)))
        return out.cpu().numpy(), valid.cpu().numpy().astype(bool)

    # ------------------------------------------------------------------ transformations
    def update_transformations(self, nnrt_data):
        """Graph-node update after estimation (warpfield.py:389-420): store R, T = -R g + g + t,
        the deformed nodes and the target frame id."""
        assert self.frame_id == self.tsdf.frame_id, \
            f"Warpfield maps to:{self.frame_id}th frame but TSDF maps to:{self.tsdf.frame_id}th frame"
        assert self.rotations.shape[0] == self.deformed_nodes.shape[0], \
            f"Nodes mismatch between transformation data:{self.rotations.shape} and deformed node:{self.deformed_nodes.shape} data"
        N = self.deformed_nodes.shape[0]
        R = nnrt_data["node_rotations"]
        # warpfield.py:407-408 loops `R[i] @ g[i]` in Python (O(M) interpreter time); the batched
        # matmul below returns the same bits (tests/test_host_logic.py pins that) in ~1/30 of the time
        self.translations = -_rot_apply(R, self.graph.nodes[:N]) + self.graph.nodes + nnrt_data["node_translations"]
        self.rotations = R
        self.deformed_nodes = nnrt_data["deformed_nodes_to_target"]
        self.frame_id = nnrt_data["target_frame_id"]
        self._transforms_dirty = True
        if self.savepath is not None:                                                  # warpfield.py:418
            np.save(os.path.join(self.savepath, "deformed_nodes", f"{self.frame_id}.npy"), self.deformed_nodes)

    def get_transformation_wrt_graph_node(self):
        """(R, t) with t relative to each node (warpfield.py:422-436)."""
        N = self.rotations.shape[0]
        translations = self.translations - self.graph.nodes[:N] + _rot_apply(self.rotations, self.graph.nodes[:N])
        return self.rotations, translations

    def get_transformation_wrt_origin(self, rotations, translations):
        """(R, T) with T relative to the origin (warpfield.py:438-449)."""
        N = rotations.shape[0]
        translations = translations + self.graph.nodes[:N] - _rot_apply(rotations, self.graph.nodes[:N])
        return rotations, translations

    def get_deformed_nodes(self):                                                      # warpfield.py:585-593
        assert self.frame_id == self.tsdf.frame_id, \
            f"Warpfield maps to:{self.frame_id}th frame but TSDF maps to:{self.tsdf.frame_id}th frame"
        assert self.graph.nodes.shape == self.translations.shape
        return self.deformed_nodes

    # ------------------------------------------------------------------ node insertion
    def find_unreachable_nodes(self, points):
        """Indices of points farther than 2*node_coverage from every node, farthest first
        (warpfield.py:455-485).  Nearest-node distance: ofx_nearest_node (exact float32 brute force on the device)."""
        pts = _capi.to_device(points, torch.float32, self.device).reshape(-1, 3)
        nd = _capi.to_device(self.graph.nodes, torch.float32, self.device)
        dist_d = torch.empty(pts.shape[0], dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _capi.check(self._lib.ofx_nearest_node(_capi.ptr(pts), pts.shape[0], _capi.ptr(nd), nd.shape[0],
                                                   _capi.ptr(dist_d), None, _capi.stream_ptr(self.device)))
        dist = dist_d.cpu().numpy()                       # float32, as the float32 KD-tree returns it (warpfield.py:463-464)
        unreachable_verts = np.where(dist > 2 * self.node_coverage)[0]
        if len(unreachable_verts) == 0:
            return []
        order = np.argsort(dist[unreachable_verts])[::-1]
        return unreachable_verts[order]

    def _nearest_rows(self, points, nodes):
        """Index of the nearest row of `nodes` for every row of `points` (float32, ties -> lower index): ofx_nearest_node."""
        pts = _capi.to_device(points, torch.float32, self.device).reshape(-1, 3)
        nd = _capi.to_device(nodes, torch.float32, self.device).reshape(-1, 3)
        idx = torch.empty(pts.shape[0], dtype=torch.int32, device=self.device)
        if pts.shape[0] == 0:
            return np.zeros(0, np.int64)
        with torch.cuda.device(self.device):
            _capi.check(self._lib.ofx_nearest_node(_capi.ptr(pts), pts.shape[0], _capi.ptr(nd), nd.shape[0], None, _capi.ptr(idx),
                                                   _capi.stream_ptr(self.device)))
        return idx.cpu().numpy().astype(np.int64)

    def update_graph(self, canonical_model=None):
        """Node insertion (warpfield.py:487-582): add nodes for canonical vertices the graph does
        not reach, re-skin the volume, initialise the new nodes' transforms.

        canonical_model: optional (vertices, faces, ...) tuple; default tsdf.get_canonical_model()
        (warpfield.py:496).  New-node transforms come from `self.model.run_arap` when a model is
        attached (warpfield.py:555); otherwise each new node copies its nearest old node's R, t."""
        self.updating_warpfield = True
        if canonical_model is None:
            canonical_model = self.tsdf.get_canonical_model()
        verts, faces = canonical_model[0], canonical_model[1]
        # warpfield.py:500-506: candidates are searched among the NON-ERODED vertices (erode_mesh, 3 requested rounds --
        # EDGraph.erode_mesh replaces any truthy count by erosion_num_iterations, as the reference does) ...
        non_eroded = self.graph.erode_mesh(verts, faces, num_iterations=3).reshape(-1) if hasattr(self.graph, "erode_mesh") \
            else np.ones(len(verts), dtype=bool)
        new_verts_indices = self.find_unreachable_nodes(verts[non_eroded])
        if len(new_verts_indices) == 0:
            self.updating_warpfield = False
            return False
        old_num_nodes = self.graph.nodes.shape[0]
        old_nodes = np.array(self.graph.nodes[:old_num_nodes], copy=True)
        R_rel, t_rel = self.get_transformation_wrt_graph_node()      # relative to the nodes as they are now
        # ... and, as in the reference (warpfield.py:516), the indices -- positions within the non-eroded subset -- are
        # handed to EDGraph.update together with the FULL vertex array.  Kept for identical results.
        update = self.graph.update(verts, faces, new_verts_indices)
        if update:
            M = self.graph.nodes.shape[0]
            kept = M >= old_num_nodes and np.array_equal(self.graph.nodes[:old_num_nodes], old_nodes)
            if not kept:
                # the edge rebuild dropped nodes with too few neighbours and renumbered the rest (edg.py:596): every node of
                # the new table takes the transform of the nearest old node (an old node that survived finds itself)
                nearest = self._nearest_rows(self.graph.nodes, old_nodes)
                R_new, t_new = np.asarray(R_rel)[nearest], np.asarray(t_rel)[nearest]
                deformed = self.graph.nodes + t_new
            elif getattr(self, "model", None) is not None and hasattr(self.graph, "get_reduced_graph"):
                valid_nodes_mask = np.zeros(M, dtype=bool)
                valid_nodes_mask[:old_num_nodes] = True
                graph_data = self.graph.get_reduced_graph(valid_nodes_mask)
                td = {"source_frame_id": self.tsdf.fopt.source_frame, "target_frame_id": self.frame_id,
                      "node_rotations": R_rel, "node_translations": t_rel,
                      "deformed_nodes_to_target": self.graph.nodes[valid_nodes_mask] + t_rel}
                est = self.model.run_arap(graph_data, td, self.graph, self)
                R_new, t_new, deformed = est["node_rotations"], est["node_translations"], est["deformed_nodes_to_target"]
            else:
                old, new = self.graph.nodes[:old_num_nodes], self.graph.nodes[old_num_nodes:]
                nearest = self._nearest_rows(new, old)
                R_new = np.concatenate([R_rel, np.asarray(R_rel)[nearest]], axis=0)
                t_new = np.concatenate([t_rel, np.asarray(t_rel)[nearest]], axis=0)
                deformed = self.graph.nodes + t_new
            self.deformed_nodes = deformed
            self.rotations, self.translations = self.get_transformation_wrt_origin(R_new, t_new)
            self._sync_device()          # new node table + anchor cache rebuild (warpfield.py:532-533)
        self.frame_id = self.tsdf.frame_id
        self.updating_warpfield = False
        return update
