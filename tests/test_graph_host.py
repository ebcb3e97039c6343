"""CPU: the host half of the graph builder in libofx.so (ofx_compute_edges_geodesic_host, ofx_node_and_edge_clean_up_host,
ofx_compute_clusters_host) through the reference's call signatures, against golden vectors produced by the reference's own
C++ (csrc/cpu/graph_proc.cpp compiled from /root/reference, oracle/make_golden_graph.py).  Edges, masks and clusters:
identical; weights and distances: bit-identical (same float32 operation order, same libm exp)."""
import numpy as np
import pytest

from conftest import load_golden


@pytest.fixture(scope="module")
def gg():
    return load_golden("ref_graph")


def _run_geo(gg, name, n_threads=0):
    from occlusionfusion_b200.graph_proc import compute_edges_geodesic_c
    E, cov, only_valid, enforce = gg[f"geo_{name}_params"]
    nidx = gg[f"geo_{name}_node_idx"]
    M, V = len(nidx), len(gg["mesh_verts"])
    e = -np.ones((M, int(E)), np.int32)
    w = np.zeros((M, int(E)), np.float32)
    d = np.zeros((M, int(E)), np.float32)
    n2v = -np.ones((M, V), np.float32)
    compute_edges_geodesic_c(gg["mesh_verts"], gg[f"geo_{name}_valid"], gg["mesh_faces"], nidx, int(E), float(cov), e, w, d, n2v,
                             bool(only_valid), bool(enforce), n_threads)
    return e, w, d, n2v


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_geodesic_edges_match_reference(gg, name):
    e, w, d, n2v = _run_geo(gg, name)
    assert np.array_equal(e, gg[f"geo_{name}_edges"])
    assert np.array_equal(d, gg[f"geo_{name}_dist"]) and np.array_equal(w, gg[f"geo_{name}_weights"])
    assert np.array_equal((n2v != -1).sum(1), gg[f"geo_{name}_n2v_count"])
    assert np.array_equal(np.where(n2v != -1, n2v, 0).astype(np.float64).sum(1), gg[f"geo_{name}_n2v_sum"])
    assert np.array_equal(n2v[:6], gg[f"geo_{name}_n2v_head"])
    e1, w1, d1, n1 = _run_geo(gg, name, n_threads=1)                    # the thread count does not change anything
    assert np.array_equal(e, e1) and np.array_equal(w, w1) and np.array_equal(d, d1) and np.array_equal(n2v, n1)
    if name == "c":
        assert (e == -1).any() and (d <= 2 * 0.05 + 1e-6).all()          # the walk stops at 2 * node_coverage


def test_clean_up_and_clusters_match_reference(gg):
    from occlusionfusion_b200.graph_proc import compute_clusters_c, node_and_edge_clean_up_c
    names = [f"geo_{n}" for n in "abc"]
    for n in names:
        e = gg[f"{n}_edges"]
        mask = np.ones((len(e), 1), bool)
        node_and_edge_clean_up_c(e, mask)
        assert np.array_equal(mask.reshape(-1), gg[f"{n}_clean"])
        cl = -np.ones((len(e), 1), np.int32)
        assert compute_clusters_c(e, cl) == gg[f"{n}_cluster_sizes"].tolist()
        assert np.array_equal(cl.reshape(-1), gg[f"{n}_clusters"])
    removed = 0
    for k in range(6):
        e = gg[f"rnd{k}_edges"]
        mask = gg[f"rnd{k}_mask_in"].reshape(-1, 1).copy()
        node_and_edge_clean_up_c(e, mask)
        assert np.array_equal(mask.reshape(-1), gg[f"rnd{k}_mask_out"]), k
        removed += int(gg[f"rnd{k}_mask_in"].sum() - mask.sum())
        cl = -np.ones((len(e), 1), np.int32)
        assert compute_clusters_c(e, cl) == gg[f"rnd{k}_sizes"].tolist(), k
        assert np.array_equal(cl.reshape(-1), gg[f"rnd{k}_clusters"]), k
    assert removed > 100


def test_graph_host_argument_errors(gg):
    from occlusionfusion_b200._capi import OfxError
    from occlusionfusion_b200.graph_proc import compute_clusters_c, compute_edges_geodesic_c
    v, f = gg["mesh_verts"], gg["mesh_faces"].copy()
    e, w, d = -np.ones((2, 4), np.int32), np.zeros((2, 4), np.float32), np.zeros((2, 4), np.float32)
    f[0, 0] = len(v) + 5
    with pytest.raises(OfxError, match="out of range"):
        compute_edges_geodesic_c(v, np.ones(len(v), bool), f, np.array([3, 9], np.int32), 4, 0.05, e, w, d, None, True, True)
    with pytest.raises(TypeError):
        compute_edges_geodesic_c(v, np.ones(len(v), bool), gg["mesh_faces"], np.array([3, 9]), 4, 0.05, e.astype(np.int64), w, d,
                                 None, True, True)
    with pytest.raises(OfxError, match="out of range"):
        compute_clusters_c(np.array([[5, -1]], np.int32), -np.ones((1, 1), np.int32))
    # a node whose vertex is an invalid (eroded) one: the reference exits the process, here it is an error
    valid = np.ones(len(v), bool)
    valid[3] = False
    with pytest.raises(OfxError, match="not a valid vertex"):
        compute_edges_geodesic_c(v, valid, gg["mesh_faces"], np.array([3, 9], np.int32), 4, 0.05, e, w, d, None, True, True)


@pytest.mark.gpu
def test_reduced_graph_clusters_and_pyramid_match_reference(gg):
    """remove_nodes_with_not_enough_neighbours / get_reduced_graph / compute_clusters / create_graph_pyramid
    (embedded_deformation_graph.py:261-480) against the reference's own Python + C++ run on the same hand-filled graph.
    (EDGraph runs these on the device: ofx_graph_edges.cu.)"""
    from occlusionfusion_b200 import EDGraph
    g = EDGraph(nodes=gg["sample_pos_1"].copy(), graph_generation_parameters={"node_coverage": float(gg["coverage"])})
    g.vertices, g.faces = gg["mesh_verts"], gg["mesh_faces"]
    g.node_indices = gg["geo_a_node_idx"].copy()
    g.edges, g.edges_weights, g.edges_distances = gg["red_edges_in"].copy(), gg["geo_a_weights"].copy(), gg["geo_a_dist"].copy()
    g.clusters = -np.ones((len(g.nodes), 1), np.int32)
    g.remove_nodes_with_not_enough_neighbours()
    assert len(g.nodes) == len(gg["red_nodes"]) < len(gg["sample_pos_1"])
    assert np.array_equal(g.nodes, gg["red_nodes"]) and np.array_equal(g.node_indices, gg["red_node_indices"])
    assert np.array_equal(g.edges, gg["red_edges"]) and np.array_equal(g.edges_distances, gg["red_dist"])
    assert np.array_equal(g.edges_weights, gg["red_weights"])
    g.compute_clusters()
    assert np.array_equal(g.clusters, gg["red_clusters"])
    g.create_graph_pyramid()
    pyd = g.get_graph_pyramid()
    keys = [k[4:] for k in gg.files if k.startswith("pyd_")]
    assert sorted(pyd) == sorted(keys) and len(keys) == 10
    for k in keys:
        assert np.array_equal(np.asarray(pyd[k]), gg["pyd_" + k]), k


def _ref_graph():
    """The reference's own C++ (oracle/_ref/ref_graph.so, built by oracle/build_ref.sh where /root/reference exists)."""
    import os
    import sys
    here = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
    if not os.path.exists(os.path.join(here, "ref_graph.so")):
        pytest.skip("oracle/_ref/ref_graph.so not built (needs /root/reference)")
    import torch  # noqa: F401
    sys.path.insert(0, here)
    try:
        import ref_graph
    except Exception as e:      # noqa: BLE001
        pytest.skip(f"ref_graph not loadable: {e}")
    return ref_graph


@pytest.mark.parametrize("jitter", [0.0, 1e-3])
def test_geodesic_edges_against_the_compiled_reference_with_ties(oracle, jitter):
    """A regular, un-jittered grid mesh makes equal path lengths the rule, so the order in which equal distances leave the heap
    decides edges and visited sets: the host builder must agree with the reference's compiled C++ there too (and on the
    jittered mesh), at a size the goldens do not carry (14 400 vertices, ~330 nodes)."""
    from occlusionfusion_b200.graph_proc import compute_clusters_c, compute_edges_geodesic_c, node_and_edge_clean_up_c
    ref = _ref_graph()
    n = 120
    rng = np.random.default_rng(4)
    u, v = np.meshgrid(np.linspace(-0.6, 0.6, n), np.linspace(-0.6, 0.6, n), indexing="ij")
    verts = (np.stack([u, v, np.ones_like(u)], -1).reshape(-1, 3) + rng.uniform(-jitter, jitter, (n * n, 3))).astype(np.float32)
    ij = np.argwhere(rng.random((n - 1, n - 1)) > 0.03)
    a = ij[:, 0] * n + ij[:, 1]
    faces = np.concatenate([np.stack([a, a + 1, a + n], 1), np.stack([a + 1, a + n + 1, a + n], 1)]).astype(np.int32)
    V = len(verts)
    ne = oracle.erode_mesh(V, faces, 1, 3)
    _, nidx = oracle.sample_nodes(verts, ne, 0.05, True)
    nidx = nidx.reshape(-1, 1)
    M = len(nidx)
    for E, cov, only_valid, enforce, valid in ((8, 0.05, True, True, ne), (4, 0.1, False, True, np.ones(V, bool)),
                                               (8, 0.05, True, False, ne)):
        out = []
        for impl in ("ours", "ref"):
            e, w, d = -np.ones((M, E), np.int32), np.zeros((M, E), np.float32), np.zeros((M, E), np.float32)
            n2v = -np.ones((M, V), np.float32)
            if impl == "ours":
                compute_edges_geodesic_c(verts, valid, faces, nidx, E, cov, e, w, d, n2v, only_valid, enforce)
            else:
                ref.compute_edges_geodesic(verts, valid.reshape(-1, 1), faces, nidx, E, cov, e, w, d, n2v, only_valid, enforce)
            out.append((e, w, d, n2v))
        for x, y in zip(*out):
            assert np.array_equal(x, y)
        e = out[0][0]
        m1, m2 = np.ones((M, 1), bool), np.ones((M, 1), bool)
        node_and_edge_clean_up_c(e, m1)
        ref.node_and_edge_clean_up(e, m2)
        assert np.array_equal(m1, m2)
        c1, c2 = -np.ones((M, 1), np.int32), -np.ones((M, 1), np.int32)
        assert compute_clusters_c(e, c1) == list(ref.compute_clusters(e, c2)) and np.array_equal(c1, c2)
    assert M > 250
