"""EDGraph -- the part of the reference's embedded deformation graph
(/root/reference/fusion_with_occlusion/embedded_deformation_graph.py) that the fusion hot path
touches: the node table, the graph parameters, node insertion (`update`, :496-542) and the node
half of graph creation (`erode_mesh` :153-172, node sampling :186-206).

These run on the device through libofx.so (ofx_erode_mesh, ofx_sample_nodes, ofx_greedy_insert:
include/ofx.h), replacing NeuralNRT._C.erode_mesh / sample_nodes (csrc/cpu/graph_proc.cpp:17-136) and
the Python loop of `update`; results are identical to the serial code, acceptance order included.
Geodesic edges, clusters and the pyramid (:207-331, graph_proc.cpp:155-481) are not rebuilt (the
fusion path reads only `nodes`); nodes may also be supplied by the caller.
"""
import ctypes
import logging

import numpy as np
import torch

from . import _capi

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
        self.log = logging.getLogger(__name__)
        self.tsdf, self.vis = tsdf, visualizer
        if tsdf is not None:
            self.opt = tsdf.fopt
            tsdf.graph = self                                   # embedded_deformation_graph.py:40
        self.graph_generation_parameters = dict(DEFAULT_GRAPH_PARAMETERS)
        if graph_generation_parameters:
            self.graph_generation_parameters.update(graph_generation_parameters)
        if nodes is None:
            raise NotImplementedError("pass nodes=, or build them with EDGraph.from_mesh(vertices, faces): marching cubes "
                                      "(tsdf.py:794) is not rebuilt, so the graph cannot start from a TSDFVolume alone")
        self.nodes = np.ascontiguousarray(nodes, dtype=np.float32)

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
        were added.  Edge / cluster / pyramid rebuild (:545-603) is a "next" row."""
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
        return True
