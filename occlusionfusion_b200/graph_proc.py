"""The host half of the reference's graph builder behind its own call signatures: numpy arrays in, outputs written in
place, as the functions of `NeuralNRT._C` that `embedded_deformation_graph.py:21-23` imports (csrc/main.cpp:30, 32, 33;
csrc/cpu/graph_proc.cpp:155-300, 388-481).  The work runs in libofx.so (include/ofx.h, `*_host` entry points: host C++,
per-node searches spread over host threads); no device is needed.
"""
import ctypes as C

import numpy as np

from . import _capi


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def _inplace(a, dtype, name):
    if not (isinstance(a, np.ndarray) and a.dtype == dtype and a.flags.c_contiguous and a.flags.writeable):
        raise TypeError(f"{name} must be a writable C-contiguous {np.dtype(dtype).name} array (it is filled in place)")
    return a


def compute_edges_geodesic_c(vertices, valid_vertices, faces, node_indices, num_neighbours, node_coverage, graph_edges,
                             graph_edges_weights, graph_edges_distances, node_to_vertex_distances,
                             allow_only_valid_vertices, enforce_total_num_neighbors, n_threads=0):
    """graph_proc.cpp:155-300.  graph_edges (M,E) i32 / weights / distances (M,E) f32 / node_to_vertex_distances (M,V) f32
    (or None) are filled in place for the neighbours found; pre-fill them as the reference's callers do (-1, 0, 0, -1)."""
    v = np.ascontiguousarray(vertices, dtype=np.float32)
    f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
    ni = np.ascontiguousarray(node_indices, dtype=np.int32).reshape(-1)
    valid = np.ascontiguousarray(np.asarray(valid_vertices).reshape(-1), dtype=np.uint8)
    M, E, V = len(ni), int(num_neighbours), len(v)
    e = _inplace(graph_edges, np.int32, "graph_edges")
    w = _inplace(graph_edges_weights, np.float32, "graph_edges_weights")
    d = _inplace(graph_edges_distances, np.float32, "graph_edges_distances")
    assert e.shape == (M, E) and w.shape == (M, E) and d.shape == (M, E) and len(valid) == V
    n2v = node_to_vertex_distances
    if n2v is not None:
        _inplace(n2v, np.float32, "node_to_vertex_distances")
        assert n2v.shape == (M, V)
    _capi.check(_capi.load().ofx_compute_edges_geodesic_host(_p(v), _p(valid), V, _p(f), len(f), _p(ni), M, E,
                                                             C.c_float(node_coverage), _p(e), _p(w), _p(d), _p(n2v),
                                                             int(bool(allow_only_valid_vertices)),
                                                             int(bool(enforce_total_num_neighbors)), int(n_threads)))


def node_and_edge_clean_up_c(graph_edges, valid_nodes_mask):
    """graph_proc.cpp:388-438.  valid_nodes_mask (M,1) bool is updated in place."""
    e = np.ascontiguousarray(graph_edges, dtype=np.int32)
    m = _inplace(valid_nodes_mask, np.bool_, "valid_nodes_mask")
    assert e.ndim == 2 and m.size == e.shape[0]
    _capi.check(_capi.load().ofx_node_and_edge_clean_up_host(_p(e), e.shape[0], e.shape[1], _p(m.view(np.uint8))))


def compute_clusters_c(graph_edges, graph_clusters):
    """graph_proc.cpp:440-481.  graph_clusters (M,1) i32 is filled in place; returns the list of cluster sizes."""
    e = np.ascontiguousarray(graph_edges, dtype=np.int32)
    c = _inplace(graph_clusters, np.int32, "graph_clusters")
    assert e.ndim == 2 and c.size == e.shape[0]
    sizes = np.zeros(max(e.shape[0], 1), np.int32)
    n = C.c_int32(0)
    _capi.check(_capi.load().ofx_compute_clusters_host(_p(e), e.shape[0], e.shape[1], _p(c), _p(sizes), C.byref(n)))
    return sizes[:n.value].tolist()


def downsample_nodes(nodes, node_coverage):
    """One pyramid level (embedded_deformation_graph.py:279-298): (down_sample_idx, up_sample_idx) as Python lists."""
    nd = np.ascontiguousarray(nodes, dtype=np.float32).reshape(-1, 3)
    M = len(nd)
    down, up = np.zeros(max(M, 1), np.int32), np.zeros(max(M, 1), np.int32)
    n = C.c_int32(0)
    _capi.check(_capi.load().ofx_downsample_nodes_host(_p(nd), M, float(node_coverage), _p(down), C.byref(n), _p(up)))
    return down[:n.value].tolist(), up[:M].tolist()


def compute_pixel_anchors_euclidean_c(graph_nodes, point_image, node_coverage, pixel_anchors, pixel_weights):
    """graph_proc.cpp:610-709 (binding csrc/main.cpp:35) on the device (ofx_pixel_anchors_euclidean): point_image (3,H,W) f32;
    pixel_anchors (i32) / pixel_weights (f32) are resized in place to (H,W,4), as the reference's callee does, and filled."""
    import torch
    _capi.require_cuda()
    dev = torch.device("cuda", torch.cuda.current_device())
    nodes = _capi.to_device(np.asarray(graph_nodes), torch.float32, dev).reshape(-1, 3)
    img = _capi.to_device(np.asarray(point_image), torch.float32, dev)
    assert img.dim() == 3 and img.shape[0] >= 3
    H, W = int(img.shape[1]), int(img.shape[2])
    a = torch.empty((H, W, 4), dtype=torch.int32, device=dev)
    w = torch.empty((H, W, 4), dtype=torch.float32, device=dev)
    _capi.check(_capi.load().ofx_pixel_anchors_euclidean(_capi.ptr(nodes) if nodes.shape[0] else None, nodes.shape[0],
                                                         _capi.ptr(img[:3].contiguous()), H, W, C.c_float(node_coverage),
                                                         _capi.ptr(a), _capi.ptr(w), _capi.stream_ptr(dev)))
    for dst, src, dt in ((pixel_anchors, a, np.int32), (pixel_weights, w, np.float32)):
        _inplace(dst, dt, "pixel_anchors / pixel_weights")
        dst.resize((H, W, 4), refcheck=False)
        dst[...] = src.cpu().numpy()
