"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/ref_graph.npz by RUNNING THE REFERENCE'S OWN CODE
in this container for the "next" rows of SURVEY.md section 8 (f2, f3, f4).
Usage:  sh oracle/build_ref.sh && python oracle/make_golden_graph.py

What runs, and from where (nothing is copied into the repo):
  EDGraph.update            the reference's Python, fusion_with_occlusion/embedded_deformation_graph.py:496-542,
                            imported in place; `self` is a small namespace whose calculate_distance_matrix
                            raises right after `self.nodes` was extended (:542), i.e. before the edge rebuild
                            (:545-603, not on these rows)
  WarpField.find_unreachable_nodes   warpfield.py:455-485, imported in place (pykdtree -> scipy shim)
  erode_mesh, sample_nodes  the reference's C++, csrc/cpu/graph_proc.cpp:17-136, compiled from /root/reference
                            into oracle/_ref/ref_graph.so (oracle/build_ref.sh, oracle/ref_graph_bind.cpp)
  compute_truncated_region  the reference's numba function, tsdf.py:706-745, imported in place
  save_graph_* / load_graph_*   utils/utils.py:205-300: the function definitions are parsed out of the
                            reference file at run time (the module itself needs matplotlib / skimage / PIL)
"""
import ast
import logging
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(HERE), "tests", "golden")


def load_ref_graph():
    import torch  # noqa: F401  (the extension header pulls in torch's pybind casters)
    sys.path.insert(0, os.path.join(HERE, "_ref"))
    import ref_graph
    return ref_graph


def import_ref_edgraph(ref_graph):
    ref_shims.install()
    c = sys.modules["NeuralNRT._C"]
    for name in ("erode_mesh", "sample_nodes"):
        setattr(c, name, getattr(ref_graph, name))
    for name in ("compute_edges_geodesic", "node_and_edge_clean_up", "compute_clusters"):
        setattr(c, name, getattr(ref_graph, name))
    c.compute_mesh_from_depth = None
    if "model" not in sys.modules:
        m = ref_shims._module("model")
        m.dataset = ref_shims._module("model.dataset")
    import embedded_deformation_graph as ref_edg
    return ref_edg


def ref_format_functions():
    """The reference's own save_graph_* / load_graph_* definitions (utils/utils.py), exec'd from the file."""
    path = os.path.join(ref_shims.REF_ROOT, "utils", "utils.py")
    tree = ast.parse(open(path).read())
    wanted = [n for n in tree.body if isinstance(n, ast.FunctionDef) and
              (n.name.startswith("save_graph_") or n.name.startswith("load_graph_"))]
    ns = {"os": os, "np": np, "struct": __import__("struct")}
    exec(compile(ast.Module(body=wanted, type_ignores=[]), path, "exec"), ns)
    return ns


class _Stop(Exception):
    pass


def grid_mesh(rng, n=56, hole_frac=0.08):
    """A depth-image-like triangulated height field with holes (what compute_mesh_from_depth yields)."""
    u, v = np.meshgrid(np.linspace(-0.5, 0.5, n), np.linspace(-0.4, 0.4, n), indexing="ij")
    z = 1.0 + 0.15 * np.sin(5 * u) * np.cos(4 * v)
    verts = np.stack([u, v, z], -1).reshape(-1, 3) + rng.uniform(-2e-3, 2e-3, (n * n, 3))
    keep = rng.random((n - 1, n - 1)) > hole_frac
    faces = []
    for i in range(n - 1):
        for j in range(n - 1):
            if keep[i, j]:
                a, b, c, d = i * n + j, i * n + j + 1, (i + 1) * n + j, (i + 1) * n + j + 1
                faces += [(a, b, c), (b, d, c)]
    return verts.astype(np.float32), np.asarray(faces, np.int32)


def adversarial_pairs(rng, coverage, want=24):
    """Vertex pairs (p, q) whose float32 squared distance is <= coverage^2 under one summation order and
    > coverage^2 under the other: they decide which order the reference's Eigen squaredNorm uses."""
    c2 = np.float32(coverage) * np.float32(coverage)
    pairs = []
    while len(pairs) < want:
        p = rng.uniform(-1, 1, (200000, 3)).astype(np.float32)
        d = rng.normal(size=(200000, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        q = (p.astype(np.float64) + d * float(coverage) * (1 + rng.uniform(-2e-7, 2e-7, (200000, 1)))).astype(np.float32)
        e = p - q
        s = e * e
        seq = (s[:, 0] + s[:, 1]) + s[:, 2]
        eig = s[:, 0] + (s[:, 1] + s[:, 2])
        idx = np.nonzero((seq <= c2) != (eig <= c2))[0]
        pairs += [(p[i], q[i]) for i in idx]
    return np.asarray(pairs[:want], np.float32)      # (want, 2, 3)


def main():
    rng = np.random.default_rng(20231)
    ref_graph = load_ref_graph()
    ref_edg = import_ref_edgraph(ref_graph)
    ref_tsdf, ref_wf = ref_shims.import_reference()
    out = {}

    # ---- f2: erode_mesh + sample_nodes (reference C++) ---------------------------------------------------------
    verts, faces = grid_mesh(rng)
    out["mesh_verts"], out["mesh_faces"] = verts, faces
    for it, mn in ((1, 3), (3, 3), (2, 4), (0, 3)):
        out[f"erode_{it}_{mn}"] = np.asarray(ref_graph.erode_mesh(verts, faces, it, mn)).reshape(-1)
    coverage = 0.05
    ne = out["erode_1_3"].reshape(-1, 1)
    for use_ne in (True, False):
        pos, idx = np.zeros(0, np.float32), np.zeros(0, np.int32)
        n = ref_graph.sample_nodes(verts, ne, pos, idx, coverage, use_ne, False)
        out[f"sample_idx_{int(use_ne)}"] = idx[:n].reshape(-1).copy()
        out[f"sample_pos_{int(use_ne)}"] = pos[:n].copy()
    pairs = adversarial_pairs(rng, coverage)
    res = []
    for p, q in pairs:
        pos, idx = np.zeros(0, np.float32), np.zeros(0, np.int32)
        res.append(ref_graph.sample_nodes(np.stack([p, q]), np.ones((2, 1), bool), pos, idx, coverage, True, False))
    out["adv_pairs"], out["adv_n_nodes"] = pairs, np.asarray(res, np.int32)
    out["coverage"] = np.float64(coverage)

    # ---- f2: geodesic edges, clean-up, clusters (reference C++) --------------------------------------------------------
    node_idx = out["sample_idx_1"].reshape(-1, 1).astype(np.int32)
    M, V = len(node_idx), len(verts)

    def ref_edges(valid, nidx, E, cov, only_valid, enforce):
        e = -np.ones((len(nidx), E), np.int32)
        w_ = np.zeros((len(nidx), E), np.float32)
        d_ = np.zeros((len(nidx), E), np.float32)
        n2v = -np.ones((len(nidx), V), np.float32)
        ref_graph.compute_edges_geodesic(verts, valid, faces, nidx, E, cov, e, w_, d_, n2v, only_valid, enforce)
        return e, w_, d_, n2v

    cases = {"a": (ne, node_idx, 8, coverage, True, True),                        # create_graph_from_mesh, edg.py:226-233
             "b": (np.ones((V, 1), bool), node_idx[::3].copy(), 4, 2 * coverage, False, True),   # pyramid level, edg.py:313-320
             "c": (ne, node_idx, 8, coverage, True, False)}                       # walk limited to 2 * coverage
    for name, (valid, nidx, E, cov, ov_, enf) in cases.items():
        e, w_, d_, n2v = ref_edges(valid, nidx, E, cov, ov_, enf)
        out[f"geo_{name}_node_idx"], out[f"geo_{name}_valid"] = nidx, valid.reshape(-1)
        out[f"geo_{name}_params"] = np.array([E, cov, ov_, enf], np.float64)
        out[f"geo_{name}_edges"], out[f"geo_{name}_weights"], out[f"geo_{name}_dist"] = e, w_, d_
        out[f"geo_{name}_n2v_count"] = (n2v != -1).sum(1).astype(np.int32)
        out[f"geo_{name}_n2v_sum"] = np.where(n2v != -1, n2v, 0).astype(np.float64).sum(1)
        out[f"geo_{name}_n2v_head"] = n2v[:6].copy()
        mask = np.ones((len(nidx), 1), bool)
        ref_graph.node_and_edge_clean_up(e, mask)
        out[f"geo_{name}_clean"] = mask.reshape(-1)
        clusters = -np.ones((len(nidx), 1), np.int32)
        sizes = ref_graph.compute_clusters(e, clusters)
        out[f"geo_{name}_clusters"], out[f"geo_{name}_cluster_sizes"] = clusters.reshape(-1), np.asarray(sizes, np.int32)
    # clean-up / clusters on random sparse edge tables (several components, nodes that lose their neighbours over several
    # passes, nodes that arrive invalid)
    for k in range(6):
        Mr, Er = 40 + 10 * k, 4
        e = -np.ones((Mr, Er), np.int32)
        for i in range(Mr):
            deg = rng.choice([0, 1, 1, 2, 2, 3, 4])
            lo, hi = (i // 14) * 14, min(Mr, (i // 14) * 14 + 14)          # components of at most 14 nodes
            nb = rng.choice([j for j in range(lo, hi) if j != i], size=min(deg, hi - lo - 1), replace=False)
            e[i, :len(nb)] = nb
        m0 = np.ones((Mr, 1), bool)
        m0[rng.choice(Mr, 3, replace=False)] = False
        m1 = m0.copy()
        ref_graph.node_and_edge_clean_up(e, m1)
        cl = -np.ones((Mr, 1), np.int32)
        sizes = ref_graph.compute_clusters(e, cl)
        out[f"rnd{k}_edges"], out[f"rnd{k}_mask_in"], out[f"rnd{k}_mask_out"] = e, m0.reshape(-1), m1.reshape(-1)
        out[f"rnd{k}_clusters"], out[f"rnd{k}_sizes"] = cl.reshape(-1), np.asarray(sizes, np.int32)

    # ---- f2: the reference's Python around them: remove_nodes_with_not_enough_neighbours / get_reduced_graph /
    #      compute_clusters / create_graph_pyramid (edg.py:261-480) on an EDGraph object filled by hand -------------------
    g_obj = object.__new__(ref_edg.EDGraph)
    g_obj.log = logging.getLogger("golden")
    g_obj.graph_generation_parameters = {"node_coverage": coverage, "min_neighbours": 2, "num_neighbours": 8}
    g_obj.vertices, g_obj.faces = verts, faces
    g_obj.nodes, g_obj.node_indices = out["sample_pos_1"].copy(), node_idx.copy()
    e_mod = out["geo_a_edges"].copy()
    for row in (5, 17, 40, 41, 120):
        e_mod[row, 1:] = -1                                   # nodes left with one neighbour -> removed, neighbours renumbered
    g_obj.edges, g_obj.edges_weights, g_obj.edges_distances = e_mod, out["geo_a_weights"].copy(), out["geo_a_dist"].copy()
    g_obj.clusters = -np.ones((M, 1), np.int32)
    out["red_edges_in"] = e_mod.copy()
    g_obj.remove_nodes_with_not_enough_neighbours()
    g_obj.compute_clusters()
    g_obj.create_graph_pyramid()
    out["red_nodes"], out["red_node_indices"] = g_obj.nodes, g_obj.node_indices
    out["red_edges"], out["red_weights"], out["red_dist"] = g_obj.edges, g_obj.edges_weights, g_obj.edges_distances
    out["red_clusters"] = g_obj.clusters
    for k, v in g_obj.pyd.items():
        out["pyd_" + k] = np.asarray(v)

    # ---- f2: find_unreachable_nodes + EDGraph.update (reference Python) -----------------------------------------
    # graph covers the left part of the mesh; the right part is "newly revealed"
    left = verts[verts[:, 0] < 0.05]
    nodes, _ = np.zeros(0, np.float32), None
    pos, idx = np.zeros(0, np.float32), np.zeros(0, np.int32)
    n = ref_graph.sample_nodes(left, np.ones((len(left), 1), bool), pos, idx, coverage, False, False)
    nodes = pos[:n].copy()
    wf_self = types.SimpleNamespace(source_frame_kdtree=ref_shims._KDTreeShim(nodes), node_coverage=coverage,
                                    log=logging.getLogger("golden"))
    dist, _ = wf_self.source_frame_kdtree.query(verts, k=1)
    cand = np.asarray(ref_wf.WarpField.find_unreachable_nodes(wf_self, verts))
    g_self = types.SimpleNamespace(nodes=nodes.copy(), graph_generation_parameters={"node_coverage": coverage},
                                   log=logging.getLogger("golden"))

    def stop(*a, **k):
        raise _Stop()
    g_self.calculate_distance_matrix = stop
    try:
        ref_edg.EDGraph.update(g_self, verts, faces, cand)
    except _Stop:
        pass
    out["upd_nodes_before"], out["upd_cand"], out["upd_nodes_after"] = nodes, cand.astype(np.int32), g_self.nodes
    out["upd_nearest_dist"] = dist.reshape(-1).astype(np.float32)
    # ---- f2: the WHOLE of EDGraph.update (edg.py:496-605: insertion, nearest vertices, edge rebuild, clean-up, clusters,
    #      pyramid) with the reference's Python and C++ ---------------------------------------------------------------------
    g_full = object.__new__(ref_edg.EDGraph)
    g_full.log = logging.getLogger("golden")
    g_full.graph_generation_parameters = {"node_coverage": coverage, "min_neighbours": 2, "num_neighbours": 8}
    g_full.nodes = nodes.copy()
    assert g_full.update(verts, faces, cand) is True
    out["full_nodes"], out["full_node_indices"] = g_full.nodes, np.asarray(g_full.node_indices)
    out["full_edges"], out["full_weights"], out["full_dist"] = g_full.edges, g_full.edges_weights, g_full.edges_distances
    out["full_clusters"] = g_full.clusters
    for k, v in g_full.pyd.items():
        out["fullpyd_" + k] = np.asarray(v)

    # second case: candidates that partly fall inside the existing graph's coverage (:522 pre-check) and a shuffled order
    cand2 = rng.permutation(len(verts))[:1500].astype(np.int32)
    g2 = types.SimpleNamespace(nodes=nodes.copy(), graph_generation_parameters={"node_coverage": coverage},
                               log=logging.getLogger("golden"), calculate_distance_matrix=stop)
    try:
        ref_edg.EDGraph.update(g2, verts, faces, cand2)
    except _Stop:
        pass
    out["upd2_cand"], out["upd2_nodes_after"] = cand2, g2.nodes

    # ---- pixel anchors (reference C++, graph_proc.cpp:610-709): a point image of the height field with holes, and a node
    #      set with exact duplicates so that the tie rule (later node first) decides ---------------------------------------
    Hh, Ww = 40, 56
    uu, vv = np.meshgrid(np.linspace(-0.5, 0.5, Ww), np.linspace(-0.4, 0.4, Hh))
    zz = 1.0 + 0.15 * np.sin(5 * uu) * np.cos(4 * vv)
    pimg = np.stack([uu, vv, zz]).astype(np.float32)
    pimg[:, rng.random((Hh, Ww)) < 0.1] = 0.0                                  # invalid pixels (z <= 0)
    pa_nodes = np.concatenate([out["sample_pos_1"][:120], out["sample_pos_1"][:15]]).astype(np.float32)
    for name, nd in (("all", pa_nodes), ("three", pa_nodes[:3])):
        pa, pw = np.zeros(0, np.int32), np.zeros(0, np.float32)
        ref_graph.compute_pixel_anchors_euclidean(nd, pimg, coverage, pa, pw)
        out[f"pix_{name}_anchors"], out[f"pix_{name}_weights"] = pa.copy(), pw.copy()
    out["pix_image"], out["pix_nodes"] = pimg, pa_nodes

    # ---- f3: compute_truncated_region (reference numba) on the reference's own integrated volume ------------------
    g = np.load(os.path.join(GOLDEN_DIR, "ref_k4.npz"))
    for key in ("tsdf_0", "tsdf_2"):
        t = np.ascontiguousarray(g[key])
        for md in (1.2, 0.5):
            m = ref_tsdf.TSDFVolume.compute_truncated_region(t, md)
            out[f"trunc_{key}_{md}"] = np.packbits(m.reshape(-1))
    # a smooth field with values inside (-0.9, 0.9) so that the neighbour test (:740) decides
    x, y, z = np.meshgrid(*(np.linspace(-1, 1, s) for s in (20, 24, 28)), indexing="ij")
    smooth = (np.sqrt(x * x + y * y + z * z) * 4 - 2).clip(-1, 1).astype(np.float32)
    smooth[rng.random(smooth.shape) < 0.02] = 1.0
    out["trunc_smooth_in"] = smooth
    for md in (1.2, 0.3):
        out[f"trunc_smooth_{md}"] = np.packbits(ref_tsdf.TSDFVolume.compute_truncated_region(smooth, md).reshape(-1))

    # ---- f4: graph .bin formats (reference utils/utils.py) -------------------------------------------------------
    fmt = ref_format_functions()
    tmp = tempfile.mkdtemp(prefix="ofx_golden_fmt_")
    f_nodes = rng.normal(size=(17, 3)).astype(np.float32)
    f_edges = rng.integers(-1, 17, (17, 8)).astype(np.int32)
    f_w = rng.random((17, 8)).astype(np.float32)
    f_def = rng.normal(size=(17, 3)).astype(np.float32)
    f_clu = rng.integers(0, 3, (17, 1)).astype(np.int32)
    arrays = {"nodes": f_nodes, "edges": f_edges, "edges_weights": f_w, "node_deformations": f_def, "clusters": f_clu}
    for kind, arr in arrays.items():
        save = fmt.get(f"save_graph_{kind}")
        if save is None:
            continue
        path = os.path.join(tmp, kind + ".bin")
        save(path, arr)
        out[f"fmt_{kind}_in"] = arr
        out[f"fmt_{kind}_bytes"] = np.frombuffer(open(path, "rb").read(), np.uint8)
        back = fmt[f"load_graph_{kind}"](path)
        assert np.array_equal(np.asarray(back).reshape(arr.shape), arr), kind

    path = os.path.join(GOLDEN_DIR, "ref_graph.npz")
    np.savez_compressed(path, **out)
    print(f"ref_graph: V={len(verts)} F={len(faces)} non-eroded(1,3)={int(out['erode_1_3'].sum())} "
          f"sampled={len(out['sample_idx_1'])}/{len(out['sample_idx_0'])} adversarial n_nodes={out['adv_n_nodes'].tolist()} "
          f"graph {len(nodes)} -> {len(g_self.nodes)} nodes ({len(cand)} candidates), case 2 -> {len(g2.nodes)} "
          f"formats={[k for k in arrays if f'fmt_{k}_bytes' in out]} -> {path} ({os.path.getsize(path)/1e3:.0f} kB)")


if __name__ == "__main__":
    main()
