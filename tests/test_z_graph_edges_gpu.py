"""GPU: the edge half of the graph builder on the device (csrc/ofx_graph_edges.cu: geodesic edges, clean-up, clusters,
pyramid down-sampling) against the golden vectors produced by the reference's own C++ (tests/golden/ref_graph.npz,
oracle/make_golden_graph.py) and against the host builder, which tests/test_graph_host.py pins to the compiled reference --
including the regular grid on which equal path lengths are the rule."""
import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gg():
    return load_golden("ref_graph")


def _device_geo(verts, valid, faces, nidx, E, cov, only_valid, enforce):
    from occlusionfusion_b200.graph_proc import DeviceMesh, compute_edges_geodesic_d
    M = len(nidx)
    e, w, d = -np.ones((M, E), np.int32), np.zeros((M, E), np.float32), np.zeros((M, E), np.float32)
    replayed = compute_edges_geodesic_d(DeviceMesh(verts, faces, valid), nidx, E, cov, e, w, d, only_valid, enforce)
    return e, w, d, replayed


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_geodesic_edges_match_reference(gg, name):
    E, cov, only_valid, enforce = gg[f"geo_{name}_params"]
    e, w, d, replayed = _device_geo(gg["mesh_verts"], gg[f"geo_{name}_valid"], gg["mesh_faces"], gg[f"geo_{name}_node_idx"], int(E),
                                    float(cov), bool(only_valid), bool(enforce))
    assert np.array_equal(e, gg[f"geo_{name}_edges"])
    assert np.array_equal(d, gg[f"geo_{name}_dist"]) and np.array_equal(w, gg[f"geo_{name}_weights"])
    if name == "c":
        assert (e == -1).any() and (d <= 2 * 0.05 + 1e-6).all()          # the walk stops at 2 * node_coverage


def test_clean_up_and_clusters_match_reference(gg):
    from occlusionfusion_b200.graph_proc import compute_clusters_d, node_and_edge_clean_up_d
    for n in [f"geo_{n}" for n in "abc"]:
        e = gg[f"{n}_edges"]
        mask = np.ones((len(e), 1), bool)
        node_and_edge_clean_up_d(e, mask)
        assert np.array_equal(mask.reshape(-1), gg[f"{n}_clean"])
        cl = -np.ones((len(e), 1), np.int32)
        assert compute_clusters_d(e, cl) == gg[f"{n}_cluster_sizes"].tolist()
        assert np.array_equal(cl.reshape(-1), gg[f"{n}_clusters"])
    removed = 0
    for k in range(6):
        e = gg[f"rnd{k}_edges"]
        mask = gg[f"rnd{k}_mask_in"].reshape(-1, 1).copy()
        node_and_edge_clean_up_d(e, mask)
        assert np.array_equal(mask.reshape(-1), gg[f"rnd{k}_mask_out"]), k
        removed += int(gg[f"rnd{k}_mask_in"].sum() - mask.sum())
        cl = -np.ones((len(e), 1), np.int32)
        assert compute_clusters_d(e, cl) == gg[f"rnd{k}_sizes"].tolist(), k
        assert np.array_equal(cl.reshape(-1), gg[f"rnd{k}_clusters"]), k
    assert removed > 100


@pytest.mark.parametrize("jitter,n", [(0.0, 120), (1e-3, 120), (1e-3, 240)])
def test_geodesic_edges_with_ties_equal_the_host_builder(oracle, jitter, n):
    """The regular grid of tests/test_graph_host.py (14 400 vertices, ~330 nodes): un-jittered, equal path lengths decide
    edges, so the flagged nodes are replayed with the reference's heap on the device; jittered, no node needs it.  The
    57 600-vertex grid (n = 240) is too large for the shared-memory search state: the global-memory variant runs."""
    from occlusionfusion_b200.graph_proc import (compute_clusters_c, compute_clusters_d, compute_edges_geodesic_c,
                                                 node_and_edge_clean_up_c, node_and_edge_clean_up_d)
    rng = np.random.default_rng(4)
    half = 0.6 * n / 120
    u, v = np.meshgrid(np.linspace(-half, half, n), np.linspace(-half, half, n), indexing="ij")
    verts = (np.stack([u, v, np.ones_like(u)], -1).reshape(-1, 3) + rng.uniform(-jitter, jitter, (n * n, 3))).astype(np.float32)
    ij = np.argwhere(rng.random((n - 1, n - 1)) > 0.03)
    a = ij[:, 0] * n + ij[:, 1]
    faces = np.concatenate([np.stack([a, a + 1, a + n], 1), np.stack([a + 1, a + n + 1, a + n], 1)]).astype(np.int32)
    V = len(verts)
    ne = oracle.erode_mesh(V, faces, 1, 3)
    _, nidx = oracle.sample_nodes(verts, ne, 0.05, True)
    nidx = nidx.reshape(-1, 1)
    M = len(nidx)
    total_replayed = 0
    for E, cov, only_valid, enforce, valid in ((8, 0.05, True, True, ne), (4, 0.1, False, True, np.ones(V, bool)),
                                               (8, 0.05, True, False, ne), (3, 0.4, False, False, np.ones(V, bool))):
        e0, w0, d0 = -np.ones((M, E), np.int32), np.zeros((M, E), np.float32), np.zeros((M, E), np.float32)
        compute_edges_geodesic_c(verts, valid, faces, nidx, E, cov, e0, w0, d0, None, only_valid, enforce)
        e, w, d, replayed = _device_geo(verts, valid, faces, nidx, E, cov, only_valid, enforce)
        total_replayed += replayed
        assert np.array_equal(e, e0) and np.array_equal(d, d0) and np.array_equal(w, w0), (E, cov, only_valid, enforce)
        m1, m2 = np.ones((M, 1), bool), np.ones((M, 1), bool)
        node_and_edge_clean_up_c(e, m1)
        node_and_edge_clean_up_d(e, m2)
        assert np.array_equal(m1, m2)
        c1, c2 = -np.ones((M, 1), np.int32), -np.ones((M, 1), np.int32)
        assert compute_clusters_c(e, c1) == compute_clusters_d(e, c2) and np.array_equal(c1, c2)
    assert M > 250
    assert (total_replayed > 50) if jitter == 0.0 else (total_replayed == 0)


def test_downsample_nodes_on_the_device_equals_the_host_loop():
    from occlusionfusion_b200.graph_proc import downsample_nodes, downsample_nodes_d
    rng = np.random.default_rng(9)
    for M, cov in ((1, 0.1), (2, 0.1), (300, 0.1), (1500, 0.07), (1500, 0.3)):
        nodes = rng.uniform(-0.5, 0.5, (M, 3)).astype(np.float32)
        nodes[M // 2:] = nodes[: M - M // 2] + np.float32(1e-4)          # near-duplicates: never kept, nearest-node ties
        d0, u0 = downsample_nodes(nodes, cov)
        d1, u1 = downsample_nodes_d(nodes, cov)
        assert d0 == d1 and u0 == u1, (M, cov)
    assert len(d0) < 200


def test_device_graph_edges_errors(gg):
    from occlusionfusion_b200._capi import OfxError
    from occlusionfusion_b200.graph_proc import compute_clusters_d
    v, f = gg["mesh_verts"], gg["mesh_faces"].copy()
    f[0, 0] = len(v) + 5
    with pytest.raises(OfxError, match="out of range"):
        _device_geo(v, np.ones(len(v), bool), f, np.array([3, 9], np.int32), 4, 0.05, True, True)
    with pytest.raises(OfxError, match="out of range"):
        compute_clusters_d(np.array([[5, -1]], np.int32), -np.ones((1, 1), np.int32))
    valid = np.ones(len(v), bool)
    valid[3] = False
    with pytest.raises(OfxError, match="not a valid vertex"):
        _device_geo(v, valid, gg["mesh_faces"], np.array([3, 9], np.int32), 4, 0.05, True, True)
    e, w, d, _ = _device_geo(v, valid, gg["mesh_faces"], np.array([-1, 9], np.int32), 4, 0.05, True, True)
    assert (e[0] == -1).all() and (e[1] >= -1).all()
