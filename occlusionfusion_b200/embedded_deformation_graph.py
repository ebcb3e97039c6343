"""EDGraph -- the part of the reference's embedded deformation graph
(/root/reference/fusion_with_occlusion/embedded_deformation_graph.py) that the fusion hot path
touches: the node table, the graph parameters, node insertion (`update`, :496-542) and the node
half of graph creation (`erode_mesh` :153-172, node sampling :186-206).

These run on the device through libofx.so (ofx_erode_mesh, ofx_sample_nodes, ofx_greedy_insert:
include/ofx.h), replacing NeuralNRT._C.erode_mesh / sample_nodes (csrc/cpu/graph_proc.cpp:17-136) and
the Python loop of `update`; results are identical to the serial code, acceptance order included.
Geodesic edges, the "not enough neighbours" clean-up, clusters and the pyramid (:207-480,
graph_proc.cpp:155-481) are host work in the reference and here: native host code in libofx.so
(graph_proc.py, `*_host` entry points) under Python mirrors of the reference's methods.  Nodes may also
be supplied by the caller (the fusion path itself reads only `nodes`).
"""
import ctypes
import json
import logging
import os

import numpy as np
import torch

from . import _capi
from .graph_proc import (DeviceMesh, compute_clusters_d, compute_edges_geodesic_batch_d, compute_edges_geodesic_d, downsample_nodes_d,
                         node_and_edge_clean_up_d)

DEFAULT_GRAPH_PARAMETERS = {           # embedded_deformation_graph.py:62-73
    "max_triangle_distance": 0.05,
    "erosion_num_iterations": 1,
    "erosion_min_neighbours": 3,
    "node_coverage": 0.05,
    "min_neighbours": 2,
    "num_neighbours": 8,
    "require_mask": True,
}


class EDGraph:
    def __init__(self, tsdf=None, visualizer=None, source_frame_data=None, *, nodes=None,
                 graph_generation_parameters=None):
        """EDGraph(tsdf, visualizer, source_frame_data) -- the reference constructor (embedded_deformation_graph.py:27-52,
        call site fusion.py:74): graph parameters from <datadir>/graph_config.json (written with the defaults when absent),
        graph built from the marching-cubes mesh of `tsdf` (create_graph_from_tsdf), results folders created.
        Keyword-only extras: nodes= supplies the node table directly (the fusion path itself reads only `nodes`),
        graph_generation_parameters= overrides entries of the configuration."""
        self.log = logging.getLogger(__name__)
        self.tsdf, self.vis = tsdf, visualizer
        if tsdf is not None:
            self.log.info(f"Creating Graph from:{tsdf.__class__.__name__}")
            self.opt = tsdf.fopt
            tsdf.graph = self                                   # embedded_deformation_graph.py:40
        if nodes is None and tsdf is not None:
            self.load_graph_config()                            # :34
        else:
            self.graph_generation_parameters = dict(DEFAULT_GRAPH_PARAMETERS)
        if graph_generation_parameters:
            self.graph_generation_parameters.update(graph_generation_parameters)
        if nodes is not None:
            self.nodes = np.ascontiguousarray(nodes, dtype=np.float32)
            return
        if tsdf is None:
            raise ValueError("EDGraph needs a TSDFVolume (the reference signature) or nodes=")
        self.create_graph_from_tsdf()                           # :43
        datadir = getattr(self.opt, "datadir", None)
        if datadir is not None:                                 # :49-52
            self.savepath = os.path.join(datadir, "results")
            os.makedirs(self.savepath, exist_ok=True)
            os.makedirs(os.path.join(self.savepath, "updated_graph"), exist_ok=True)

    def load_graph_config(self):
        """Parameters for generating the graph (embedded_deformation_graph.py:54-75): <datadir>/graph_config.json when it
        exists, else the defaults, which are then written there."""
        datadir = getattr(self.opt, "datadir", None)
        path = os.path.join(datadir, "graph_config.json") if datadir is not None else None
        if path is not None and os.path.isfile(path):
            with open(path) as f:
                self.graph_generation_parameters = json.load(f)
        else:
            self.graph_generation_parameters = dict(DEFAULT_GRAPH_PARAMETERS)
            if path is not None:
                with open(path, "w") as f:
                    json.dump(self.graph_generation_parameters, f)

    def create_graph_from_tsdf(self):
        """embedded_deformation_graph.py:79-84: the graph of the volume's current surface (marching cubes on the device)."""
        assert hasattr(self, "tsdf"), "TSDF not defined in graph. Add tsdf as attribute to EDGraph first."
        vertices, faces, normals, colors = self.tsdf.get_mesh()
        self.log.debug(f"Creating graph from mesh. Vertices:{vertices.shape} Faces:{faces.shape}")
        self.create_graph_from_mesh(vertices, faces)

    # ------------------------------------------------------------------ device helpers
    @property
    def device(self):
        if getattr(self, "tsdf", None) is not None and hasattr(self.tsdf, "device"):
            return self.tsdf.device
        return torch.device("cuda", torch.cuda.current_device())

    @classmethod
    def from_mesh(cls, vertices, faces, tsdf=None, visualizer=None, graph_generation_parameters=None):
        """The node half of create_graph_from_mesh (embedded_deformation_graph.py:174-206): erode the mesh, then sample
        nodes node_coverage apart from the non-eroded vertices in index order.  Sets `nodes` and `node_indices`."""
        g = cls(tsdf, visualizer, nodes=np.zeros((0, 3), np.float32), graph_generation_parameters=graph_generation_parameters)
        assert vertices.shape[0] > 0 and faces.shape[0] > 0                                   # :184
        non_eroded = g.erode_mesh(vertices, faces)
        g.vertices, g.faces = vertices, faces
        g.nodes, g.node_indices = g.sample_nodes(vertices, non_eroded)
        return g

    def create_graph_from_mesh(self, vertices, faces):
        """The whole of embedded_deformation_graph.py:174-256: erode, sample nodes (device), geodesic edges between them
        (host threads), drop nodes with too few neighbours, clusters, pyramid.  Sets nodes, node_indices, edges,
        edges_weights, edges_distances, clusters, pyd, vertices, faces."""
        num_neighbours = self.graph_generation_parameters["num_neighbours"]
        node_coverage = self.graph_generation_parameters["node_coverage"]
        assert vertices.shape[0] > 0 and faces.shape[0] > 0                                   # :184
        valid_vertices = self.erode_mesh(vertices, faces)
        node_coords, node_indices = self.sample_nodes(vertices, valid_vertices)
        num_nodes = node_coords.shape[0]
        graph_edges = -np.ones((num_nodes, num_neighbours), dtype=np.int32)                   # :217-220
        graph_edges_weights = np.zeros((num_nodes, num_neighbours), dtype=np.float32)
        graph_edges_distances = np.zeros((num_nodes, num_neighbours), dtype=np.float32)
        # node_to_vertex_distances (num_nodes x num_vertices) is only scratch in the reference (:221, 247): not materialised
        self._mesh = DeviceMesh(vertices, faces, np.ones_like(valid_vertices), self.device)   # uploaded once, reused by the pyramid
        compute_edges_geodesic_d(self._mesh, node_indices, num_neighbours, node_coverage, graph_edges, graph_edges_weights,
                                 graph_edges_distances, True, True)                           # :224-233
        self.vertices, self.faces = vertices, faces
        self.nodes, self.node_indices = node_coords, node_indices
        self.edges, self.edges_weights, self.edges_distances = graph_edges, graph_edges_weights, graph_edges_distances
        self.clusters = -np.ones((num_nodes, 1), dtype=np.int32)
        self.remove_nodes_with_not_enough_neighbours()
        self.compute_clusters()
        self.create_graph_pyramid()

    def remove_nodes_with_not_enough_neighbours(self):
        """embedded_deformation_graph.py:334-368: nodes left with at most one neighbour go (repeatedly), the others are
        renumbered."""
        valid_nodes_mask = np.ones((self.nodes.shape[0], 1), dtype=bool)
        node_and_edge_clean_up_d(self.edges, valid_nodes_mask, self.device)
        reduced = self.get_reduced_graph(valid_nodes_mask)
        self.nodes = reduced["valid_nodes_at_source"]
        self.edges = reduced["graph_edges"]
        self.edges_weights = reduced["graph_edges_weights"]
        self.edges_distances = reduced["graph_edges_distances"]
        self.clusters = reduced["graph_clusters"]
        self.node_indices = self.node_indices[reduced["valid_nodes_mask"]]                   # :368 (flattens to 1-D, as there)

    def compute_clusters(self):
        """Connected components of the node graph into self.clusters (embedded_deformation_graph.py:371-380)."""
        clusters_size_list = compute_clusters_d(self.edges, self.clusters, self.device)
        for i, cluster_size in enumerate(clusters_size_list):
            if cluster_size <= 2:
                self.log.error(f"Cluster is too small {clusters_size_list}")
                self.log.error(f"It only has nodes:{np.where(self.clusters == i)[0]}")
        return clusters_size_list

    def get_reduced_graph(self, valid_nodes_mask):
        """The graph restricted to the nodes with valid_nodes_mask set (embedded_deformation_graph.py:382-477): rows of the
        dropped nodes go, neighbour ids are renumbered, dropped neighbours are squeezed out of each row (order kept) and the
        row's weights renormalised."""
        keep = np.asarray(valid_nodes_mask).squeeze()
        black = np.where(np.asarray(valid_nodes_mask) == False)[0].tolist()                   # noqa: E712
        node_coords = self.nodes[keep]
        graph_edges = self.edges[keep]
        graph_edges_weights = self.edges_weights[keep]
        graph_edges_distances = self.edges_distances[keep]
        graph_clusters = self.clusters[keep]
        num_nodes = node_coords.shape[0]
        if num_nodes == 0:
            raise RuntimeError("No nodes left in the graph")                                  # the reference exit()s, :400-402
        if len(black) > 0:
            new_id = np.where(keep, np.cumsum(keep) - 1, -1).astype(np.int32)                 # old id -> new id, -1 = dropped
            ok = ~np.isin(graph_edges, black)                                                 # -1 entries count as kept (:424)
            order = np.argsort(~ok, axis=1, kind="stable")                                    # kept entries first, order kept
            n_ok = ok.sum(axis=1, keepdims=True)
            front = np.arange(graph_edges.shape[1])[None, :] < n_ok
            ids = np.take_along_axis(graph_edges, order, axis=1)
            graph_edges = np.where(front, np.where(ids == -1, -1, new_id[np.maximum(ids, 0)]), -1).astype(np.int32)
            graph_edges_weights = np.where(front, np.take_along_axis(graph_edges_weights, order, axis=1), 0).astype(np.float32)
            graph_edges_distances = np.where(front, np.take_along_axis(graph_edges_distances, order, axis=1), 0).astype(np.float32)
            sum_weights = np.sum(graph_edges_weights, axis=1, keepdims=True)
            graph_edges_weights = np.where(sum_weights > 0, graph_edges_weights / (sum_weights + 1e-6),
                                           graph_edges_weights).astype(np.float32)             # :452-454
        return {
            "all_nodes_at_source": self.nodes.copy(), "valid_nodes_at_source": node_coords, "graph_edges": graph_edges,
            "graph_edges_weights": graph_edges_weights, "graph_edges_distances": graph_edges_distances,
            "graph_clusters": graph_clusters, "num_nodes": np.array(num_nodes, dtype=np.int64),
            "valid_nodes_mask": valid_nodes_mask,
        }

    def create_graph_pyramid(self):
        """Four-level node pyramid for motion completion (embedded_deformation_graph.py:261-331): level l keeps, in node
        order, the nodes at least 2^l * node_coverage away from every node kept before them; up_sample_idx maps a node of
        the finer level to the nearest node kept before it; each level gets geodesic edges (8, 6, 4, 3 neighbours)."""
        node_coverage = self.graph_generation_parameters["node_coverage"]
        num_neighbours_list = [8, 6, 4, 3]
        pyd = {"nn_index_l0": self.edges}
        old_nodes, level_node_indices = self.nodes, self.node_indices
        mesh = getattr(self, "_mesh", None)
        if mesh is None or mesh.V != self.vertices.shape[0] or mesh.F != np.shape(self.faces)[0]:
            mesh = self._mesh = DeviceMesh(self.vertices, self.faces, np.ones(self.vertices.shape[0], dtype=bool), self.device)
        # the node sets of the three coarser levels first (pure geometry), then ONE launch for their geodesic edges
        jobs = []
        for level in range(1, 4):
            node_coverage = node_coverage * 2
            down_sample_idx, up_sample_idx = downsample_nodes_d(old_nodes, node_coverage, self.device)   # :279-298, on the device
            E = num_neighbours_list[level]
            level_node_indices = level_node_indices[down_sample_idx]
            n = level_node_indices.shape[0]
            graph_edges = -np.ones((n, E), dtype=np.int32)
            graph_edges_weights = np.zeros((n, E), dtype=np.float32)
            graph_edges_distances = np.zeros((n, E), dtype=np.float32)
            jobs.append((level_node_indices, E, node_coverage, graph_edges, graph_edges_weights, graph_edges_distances,
                         False, True))                                                        # :313-320
            pyd["down_sample_idx" + str(level)] = down_sample_idx
            pyd["up_sample_idx" + str(level)] = up_sample_idx
            pyd["nn_index_l" + str(level)] = graph_edges
            old_nodes = old_nodes[down_sample_idx]
        compute_edges_geodesic_batch_d(mesh, jobs)
        self.pyd = pyd

    def get_graph_pyramid(self):
        return self.pyd

    def erode_mesh(self, vertices, faces, num_iterations=-1):
        """(V,1) bool, True = vertex survives the erosion (embedded_deformation_graph.py:153-172).  As in the reference,
        any truthy num_iterations (the default -1 included) is replaced by erosion_num_iterations (:166-167)."""
        _capi.require_cuda()
        if num_iterations:
            num_iterations = self.graph_generation_parameters["erosion_num_iterations"]
        V = int(np.shape(vertices)[0])
        dev = self.device
        f = _capi.to_device(np.asarray(faces), torch.int32, dev).reshape(-1, 3)
        out = torch.empty(V, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _capi.check(_capi.load().ofx_erode_mesh(_capi.ptr(f), f.shape[0], V, int(num_iterations),
                                                    int(self.graph_generation_parameters["erosion_min_neighbours"]),
                                                    _capi.ptr(out), _capi.stream_ptr(dev)))
        return out.cpu().numpy().astype(bool).reshape(-1, 1)

    def sample_nodes(self, vertices, non_eroded_vertices, use_only_valid_vertices=True):
        """(node_coords (n,3) f32, node_indices (n,1) i32): sample_nodes_c with SAMPLE_RANDOM_SHUFFLE = False
        (embedded_deformation_graph.py:191-209)."""
        _capi.require_cuda()
        dev = self.device
        v = _capi.to_device(np.asarray(vertices), torch.float32, dev).reshape(-1, 3)
        ne = _capi.to_device(np.asarray(non_eroded_vertices).reshape(-1).astype(np.uint8), torch.uint8, dev)
        V = v.shape[0]
        idx = torch.empty(max(V, 1), dtype=torch.int32, device=dev)
        pos = torch.empty((max(V, 1), 3), dtype=torch.float32, device=dev)
        n = torch.zeros(1, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            _capi.check(_capi.load().ofx_sample_nodes(_capi.ptr(v), V, _capi.ptr(ne),
                                                      ctypes.c_float(self.graph_generation_parameters["node_coverage"]),
                                                      int(bool(use_only_valid_vertices)), _capi.ptr(idx), _capi.ptr(pos),
                                                      _capi.ptr(n), _capi.stream_ptr(dev)))
        k = int(n.item())
        return pos[:k].cpu().numpy(), idx[:k].cpu().numpy().reshape(-1, 1)

    def update(self, canonical_model_vertices, canonical_model_faces, new_verts_indices, plot_update=False):
        """Greedy node insertion (embedded_deformation_graph.py:496-542): walk the candidate
        vertices in the given order (farthest first, warpfield.py:469-483); accept a vertex when it
        is at least node_coverage away from every existing node and every accepted vertex; append
        accepted vertices to `nodes` (old nodes keep their indices, :542).  Returns True when nodes
        were added.  With faces given, the edges / clusters / pyramid are rebuilt on the new mesh as the reference does
        (:545-603, `rebuild_edges`); faces=None appends the nodes only."""
        if len(new_verts_indices) == 0:
            return False
        _capi.require_cuda()
        node_coverage = self.graph_generation_parameters["node_coverage"]
        dev = self.device
        verts = _capi.to_device(np.asarray(canonical_model_vertices), torch.float32, dev).reshape(-1, 3)
        cand = _capi.to_device(np.asarray(new_verts_indices).astype(np.int32), torch.int32, dev)
        nodes = _capi.to_device(self.nodes, torch.float32, dev).reshape(-1, 3)
        accepted = torch.empty(cand.shape[0], dtype=torch.int32, device=dev)
        n = torch.zeros(1, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            _capi.check(_capi.load().ofx_greedy_insert(_capi.ptr(verts), verts.shape[0], _capi.ptr(cand), cand.shape[0],
                                                       _capi.ptr(nodes) if nodes.shape[0] else None, nodes.shape[0],
                                                       float(node_coverage), _capi.ptr(accepted), _capi.ptr(n),
                                                       _capi.stream_ptr(dev)))
        k = int(n.item())
        if k == 0:                                                                             # :538-539
            return False
        new_nodes = np.asarray(canonical_model_vertices)[accepted[:k].cpu().numpy()]
        self.nodes = np.concatenate([self.nodes, new_nodes], axis=0).astype(np.float32)          # :542
        if canonical_model_faces is not None:
            self.rebuild_edges(canonical_model_vertices, canonical_model_faces)                  # :545-603
        return True

    def rebuild_edges(self, canonical_model_vertices, canonical_model_faces):
        """The second half of the reference's update (embedded_deformation_graph.py:545-603): anchor every node on its
        nearest vertex of the new canonical mesh (ofx_argmin_norm on the device instead of the dense nodes x vertices
        matrix), geodesic edges between the nodes, drop nodes with too few neighbours, clusters, pyramid."""
        _capi.require_cuda()
        dev = self.device
        verts_np = np.ascontiguousarray(canonical_model_vertices, dtype=np.float32)
        verts = _capi.to_device(verts_np, torch.float32, dev).reshape(-1, 3)
        nodes = _capi.to_device(self.nodes, torch.float32, dev).reshape(-1, 3)
        idx = torch.empty(nodes.shape[0], dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            _capi.check(_capi.load().ofx_argmin_norm(_capi.ptr(nodes), nodes.shape[0], _capi.ptr(verts), verts.shape[0], None,
                                                     _capi.ptr(idx), _capi.stream_ptr(dev)))
        self.node_indices = idx.cpu().numpy().astype(np.int64)[:, None]                          # :547-550
        num_nodes = self.nodes.shape[0]
        num_neighbours = self.graph_generation_parameters["num_neighbours"]
        node_coverage = self.graph_generation_parameters["node_coverage"]
        graph_edges = -np.ones((num_nodes, num_neighbours), dtype=np.int32)
        graph_edges_weights = np.zeros((num_nodes, num_neighbours), dtype=np.float32)
        graph_edges_distances = np.zeros((num_nodes, num_neighbours), dtype=np.float32)
        self._mesh = DeviceMesh(verts_np, canonical_model_faces, np.ones(verts_np.shape[0], dtype=bool), dev)
        compute_edges_geodesic_d(self._mesh, self.node_indices, num_neighbours, node_coverage, graph_edges, graph_edges_weights,
                                 graph_edges_distances, True, True)                              # :575-582
        self.vertices, self.faces = canonical_model_vertices, canonical_model_faces
        self.edges, self.edges_weights, self.edges_distances = graph_edges, graph_edges_weights, graph_edges_distances
        self.clusters = -np.ones((num_nodes, 1), dtype=np.int32)
        self.remove_nodes_with_not_enough_neighbours()
        self.compute_clusters()
        self.create_graph_pyramid()
