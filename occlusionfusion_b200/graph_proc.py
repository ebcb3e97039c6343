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


# ---- the same four on the DEVICE (ofx_graph_edges.cu): what EDGraph uses ---------------------------------------------------
class DeviceMesh:
    """A mesh's vertices / faces (and optional vertex mask) uploaded once and reused by the edge builder of every pyramid
    level (embedded_deformation_graph.py:226-233, 313-320: four searches over the same mesh)."""

    def __init__(self, vertices, faces, valid_vertices=None, device=None):
        import torch
        _capi.require_cuda()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.verts = _capi.to_device(np.ascontiguousarray(vertices, dtype=np.float32), torch.float32, self.device).reshape(-1, 3)
        self.faces = _capi.to_device(np.ascontiguousarray(faces, dtype=np.int32), torch.int32, self.device).reshape(-1, 3)
        self.valid = None if valid_vertices is None else \
            _capi.to_device(np.ascontiguousarray(np.asarray(valid_vertices).reshape(-1), dtype=np.uint8), torch.uint8, self.device)
        self.V, self.F = int(self.verts.shape[0]), int(self.faces.shape[0])


def compute_edges_geodesic_batch_d(mesh, jobs):
    """Several compute_edges_geodesic searches (graph_proc.cpp:155-300) over ONE mesh in one launch.  jobs: list of tuples
    (node_indices, num_neighbours, node_coverage, graph_edges, graph_edges_weights, graph_edges_distances,
    allow_only_valid_vertices, enforce_total_num_neighbors); each job's (M,E) arrays are filled in place for the neighbours
    found (pre-filled by the caller, as in the reference).  Returns the number of nodes whose answer depended on equal
    distances (replayed with the reference's heap)."""
    import torch
    nis, counts, Es, covs, flags = [], [], [], [], []
    for ni, E, cov, e, w, d, only_valid, enforce in jobs:
        ni = np.ascontiguousarray(ni, dtype=np.int32).reshape(-1)
        M, E = len(ni), int(E)
        _inplace(e, np.int32, "graph_edges"); _inplace(w, np.float32, "graph_edges_weights"); _inplace(d, np.float32, "graph_edges_distances")
        assert e.shape == (M, E) and w.shape == (M, E) and d.shape == (M, E)
        if ni.max(initial=-1) >= mesh.V:
            raise _capi.OfxError(f"compute_edges_geodesic: node vertex {int(ni.max())} out of range")
        if only_valid and mesh.valid is None:
            raise _capi.OfxError("compute_edges_geodesic: allow_only_valid_vertices needs the vertex mask")
        nis.append(ni); counts.append(M); Es.append(E); covs.append(float(cov))
        flags.append((1 if only_valid else 0) | (2 if enforce else 0))
    total = sum(counts)
    if total == 0:
        return 0
    dev = mesh.device
    n_jobs = len(jobs)
    i32a = lambda v: (C.c_int32 * n_jobs)(*v)                                  # noqa: E731
    with torch.cuda.device(dev):
        ni_d = torch.from_numpy(np.concatenate(nis)).to(dev)
        e_d = torch.from_numpy(np.concatenate([j[3].reshape(-1) for j in jobs])).to(dev)
        d_d = torch.from_numpy(np.concatenate([j[5].reshape(-1) for j in jobs])).to(dev)
        n_d = torch.zeros(total, dtype=torch.int32, device=dev)
        status = (C.c_int32 * 2)()
        _capi.check(_capi.load().ofx_compute_edges_geodesic_batch(
            _capi.ptr(mesh.verts), _capi.ptr(mesh.valid), mesh.V, _capi.ptr(mesh.faces), mesh.F, _capi.ptr(ni_d), n_jobs, i32a(counts),
            i32a(Es), (C.c_float * n_jobs)(*covs), i32a(flags), _capi.ptr(e_d), _capi.ptr(d_d), _capi.ptr(n_d), status,
            _capi.stream_ptr(dev)))
        e_all, d_all, found = e_d.cpu().numpy(), d_d.cpu().numpy(), n_d.cpu().numpy()
    o = n0 = 0
    lib = _capi.load()
    for (ni, E, cov, e, w, d, _, _), M in zip(jobs, counts):
        E = int(E)
        e[...] = e_all[o:o + M * E].reshape(M, E)
        d[...] = d_all[o:o + M * E].reshape(M, E)
        if M:
            _capi.check(lib.ofx_edge_weights_host(_p(d), _p(np.ascontiguousarray(found[n0:n0 + M])), M, E, C.c_float(cov), _p(w)))
        o += M * E
        n0 += M
    return int(status[1])


def compute_edges_geodesic_d(mesh, node_indices, num_neighbours, node_coverage, graph_edges, graph_edges_weights,
                             graph_edges_distances, allow_only_valid_vertices, enforce_total_num_neighbors):
    """compute_edges_geodesic (graph_proc.cpp:155-300) on the device: `mesh` is a DeviceMesh; graph_edges (M,E) i32 /
    weights / distances (M,E) f32 are filled in place for the neighbours found (pre-filled by the caller, as in the reference).
    Returns the number of nodes whose answer depended on equal distances (replayed with the reference's heap)."""
    return compute_edges_geodesic_batch_d(mesh, [(node_indices, num_neighbours, node_coverage, graph_edges, graph_edges_weights,
                                                  graph_edges_distances, allow_only_valid_vertices, enforce_total_num_neighbors)])


def node_and_edge_clean_up_d(graph_edges, valid_nodes_mask, device=None):
    """node_and_edge_clean_up (graph_proc.cpp:388-438) on the device; valid_nodes_mask (M,1) bool is updated in place."""
    import torch
    _capi.require_cuda()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    e = np.ascontiguousarray(graph_edges, dtype=np.int32)
    m = _inplace(valid_nodes_mask, np.bool_, "valid_nodes_mask")
    assert e.ndim == 2 and m.size == e.shape[0]
    if e.shape[0] == 0:
        return
    with torch.cuda.device(dev):
        e_d = torch.from_numpy(e).to(dev)
        m_d = torch.from_numpy(m.view(np.uint8).reshape(-1).copy()).to(dev)
        _capi.check(_capi.load().ofx_node_and_edge_clean_up(_capi.ptr(e_d), e.shape[0], e.shape[1], _capi.ptr(m_d),
                                                            _capi.stream_ptr(dev)))
        m.reshape(-1)[...] = m_d.cpu().numpy().astype(bool)


def compute_clusters_d(graph_edges, graph_clusters, device=None):
    """compute_clusters (graph_proc.cpp:440-481) on the device; graph_clusters (M,1) i32 is filled in place; returns the
    list of cluster sizes."""
    import torch
    _capi.require_cuda()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    e = np.ascontiguousarray(graph_edges, dtype=np.int32)
    c = _inplace(graph_clusters, np.int32, "graph_clusters")
    assert e.ndim == 2 and c.size == e.shape[0]
    M = e.shape[0]
    if M == 0:
        return []
    with torch.cuda.device(dev):
        e_d = torch.from_numpy(e).to(dev)
        c_d = torch.empty(M, dtype=torch.int32, device=dev)
        s_d = torch.empty(M, dtype=torch.int32, device=dev)
        n_d = torch.zeros(1, dtype=torch.int32, device=dev)
        _capi.check(_capi.load().ofx_compute_clusters(_capi.ptr(e_d), M, e.shape[1], _capi.ptr(c_d), _capi.ptr(s_d), _capi.ptr(n_d),
                                                      _capi.stream_ptr(dev)))
        c.reshape(-1)[...] = c_d.cpu().numpy()
        return s_d[: int(n_d.item())].cpu().numpy().tolist()


def downsample_nodes_d(nodes, node_coverage, device=None):
    """One pyramid level (embedded_deformation_graph.py:279-298) on the device: the greedy rule is ofx_greedy_insert over the
    nodes in index order (no existing nodes), up_sample_idx is ofx_upsample_index.  Returns (down_sample_idx, up_sample_idx)
    as Python lists."""
    import torch
    _capi.require_cuda()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    nd = np.ascontiguousarray(nodes, dtype=np.float32).reshape(-1, 3)
    M = len(nd)
    if M == 0:
        return [], []
    lib = _capi.load()
    with torch.cuda.device(dev):
        nd_d = torch.from_numpy(nd).to(dev)
        cand = torch.arange(M, dtype=torch.int32, device=dev)
        acc = torch.empty(M, dtype=torch.int32, device=dev)
        n_d = torch.zeros(1, dtype=torch.int32, device=dev)
        S = _capi.stream_ptr(dev)
        _capi.check(lib.ofx_greedy_insert(_capi.ptr(nd_d), M, _capi.ptr(cand), M, None, 0, float(node_coverage), _capi.ptr(acc),
                                          _capi.ptr(n_d), S))
        k = int(n_d.item())
        up = torch.empty(M, dtype=torch.int32, device=dev)
        _capi.check(lib.ofx_upsample_index(_capi.ptr(nd_d), M, _capi.ptr(acc), k, _capi.ptr(up), S))
        return acc[:k].cpu().numpy().tolist(), up.cpu().numpy().tolist()
