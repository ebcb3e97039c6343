"""EDGraph -- the part of the reference's embedded deformation graph
(/root/reference/fusion_with_occlusion/embedded_deformation_graph.py) that the fusion hot path
touches: the node table, the graph parameters, and node insertion (`update`, :496-609).

Building the graph from a TSDF mesh (erosion, geodesic edges, clusters, pyramid; :174-331 and
csrc/cpu/graph_proc.cpp) is a "next" row and is not rebuilt here; nodes are
supplied by the caller.
"""
import logging

import numpy as np

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
            raise NotImplementedError("graph creation from the TSDF mesh is a 'next' row (SURVEY.md 8 f2); pass nodes=")
        self.nodes = np.ascontiguousarray(nodes, dtype=np.float32)

    def update(self, canonical_model_vertices, canonical_model_faces, new_verts_indices, plot_update=False):
        """Greedy node insertion (embedded_deformation_graph.py:496-542): walk the candidate
        vertices in the given order (farthest first, warpfield.py:469-483); accept a vertex when it
        is at least node_coverage away from every existing node and every accepted vertex; append
        accepted vertices to `nodes` (old nodes keep their indices, :542).  Returns True when nodes
        were added.  Edge / cluster / pyramid rebuild (:545-603) is a "next" row."""
        if len(new_verts_indices) == 0:
            return False
        node_coverage = self.graph_generation_parameters["node_coverage"]
        verts = canonical_model_vertices
        new_node_indices = []
        for x in new_verts_indices:
            if np.min(np.linalg.norm(self.nodes - verts[x], axis=1)) < node_coverage:
                continue
            if len(new_node_indices) == 0:
                new_node_indices.append(x)
                continue
            if np.min(np.linalg.norm(verts[new_node_indices] - verts[x], axis=1)) < node_coverage:
                continue
            new_node_indices.append(x)
        if len(new_node_indices) == 0:
            return False
        self.nodes = np.concatenate([self.nodes, verts[new_node_indices]], axis=0).astype(np.float32)
        return True
