"""GPU parity for the "next" rows: node-insertion machinery and the truncated-region
mask, through the drop-in classes (C ABI of libofx.so), against (1) golden vectors produced by RUNNING THE REFERENCE
(its Python EDGraph.update / numba compute_truncated_region and its C++ erode_mesh / sample_nodes compiled from
/root/reference; oracle/make_golden_graph.py) and (2) the C oracle on larger seeded inputs.  Index / mask outputs:
bit-exact, acceptance order included."""
import types

import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gg():
    return load_golden("ref_graph")


def _graph(nodes, coverage=0.05, **params):
    from occlusionfusion_b200 import EDGraph
    return EDGraph(nodes=nodes, graph_generation_parameters={"node_coverage": float(coverage), **params})


def _big_mesh(rng, n=300, hole_frac=0.05):
    u, v = np.meshgrid(np.linspace(-0.8, 0.8, n), np.linspace(-0.6, 0.6, n), indexing="ij")
    z = 1.0 + 0.2 * np.sin(4 * u) * np.cos(3 * v)
    verts = (np.stack([u, v, z], -1).reshape(-1, 3) + rng.uniform(-1e-3, 1e-3, (n * n, 3))).astype(np.float32)
    ij = np.argwhere(rng.random((n - 1, n - 1)) > hole_frac)
    a = ij[:, 0] * n + ij[:, 1]
    faces = np.concatenate([np.stack([a, a + 1, a + n], 1), np.stack([a + 1, a + n + 1, a + n], 1)]).astype(np.int32)
    return verts, faces


def test_erode_mesh_golden_and_oracle(gg, oracle):
    V = len(gg["mesh_verts"])
    for it, mn in ((1, 3), (3, 3), (2, 4)):
        g = _graph(np.zeros((1, 3), np.float32), erosion_num_iterations=it, erosion_min_neighbours=mn)
        got = g.erode_mesh(gg["mesh_verts"], gg["mesh_faces"])
        assert got.shape == (V, 1) and got.dtype == bool
        assert np.array_equal(got.reshape(-1), gg[f"erode_{it}_{mn}"]), (it, mn)
    # num_iterations = 0 is falsy -> no erosion rounds (embedded_deformation_graph.py:166)
    got = _graph(np.zeros((1, 3), np.float32)).erode_mesh(gg["mesh_verts"], gg["mesh_faces"], num_iterations=0)
    assert np.array_equal(got.reshape(-1), gg["erode_0_3"])
    verts, faces = _big_mesh(np.random.default_rng(5))
    for it in (1, 4):
        g = _graph(np.zeros((1, 3), np.float32), erosion_num_iterations=it)
        assert np.array_equal(g.erode_mesh(verts, faces).reshape(-1), oracle.erode_mesh(len(verts), faces, it, 3))


def test_sample_nodes_golden_and_oracle(gg, oracle):
    g = _graph(np.zeros((1, 3), np.float32), float(gg["coverage"]))
    for use_ne in (1, 0):
        pos, idx = g.sample_nodes(gg["mesh_verts"], gg["erode_1_3"], bool(use_ne))
        assert idx.shape[1] == 1 and idx.dtype == np.int32 and pos.dtype == np.float32
        assert np.array_equal(idx.reshape(-1), gg[f"sample_idx_{use_ne}"])
        assert np.array_equal(pos, gg[f"sample_pos_{use_ne}"])
    # pairs that pin the float32 summation order of the reference's compiled Eigen squaredNorm
    for (p, q), n in zip(gg["adv_pairs"], gg["adv_n_nodes"]):
        pos, idx = g.sample_nodes(np.stack([p, q]), np.ones(2, bool))
        assert len(idx) == n
    # 90 000 vertices, ~700 nodes, against the oracle; and EDGraph.from_mesh = erode + sample
    from occlusionfusion_b200 import EDGraph
    verts, faces = _big_mesh(np.random.default_rng(6))
    ne = oracle.erode_mesh(len(verts), faces, 1, 3)
    opos, oidx = oracle.sample_nodes(verts, ne, 0.05, True)
    built = EDGraph.from_mesh(verts, faces)
    assert len(oidx) > 300
    assert np.array_equal(built.node_indices.reshape(-1), oidx) and np.array_equal(built.nodes, opos)
    # nothing to sample from
    pos, idx = g.sample_nodes(verts[:64], np.zeros(64, bool))
    assert pos.shape == (0, 3) and idx.shape == (0, 1)


def test_greedy_update_golden_and_oracle(gg, oracle):
    before = gg["upd_nodes_before"]
    for cand, after in ((gg["upd_cand"], gg["upd_nodes_after"]), (gg["upd2_cand"], gg["upd2_nodes_after"])):
        g = _graph(before.copy(), float(gg["coverage"]))
        assert g.update(gg["mesh_verts"], gg["mesh_faces"], cand) is True
        assert g.nodes.dtype == np.float32 and np.array_equal(g.nodes, after)
    # every candidate inside the existing coverage -> no update (embedded_deformation_graph.py:522, 538-539)
    g = _graph(before.copy(), float(gg["coverage"]))
    near = np.nonzero(oracle.nearest_node(gg["mesh_verts"], before)[0] < 0.02)[0][:50]
    assert g.update(gg["mesh_verts"], None, near) is False and len(g.nodes) == len(before)
    # small analytic case: 1 accepted, 2 too close to 1, 3 accepted, 4 too close to node 1
    g = _graph(np.array([[0, 0, 0], [1, 0, 0]], np.float32), 0.5)
    verts = np.array([[0.1, 0, 0], [2, 0, 0], [2.2, 0, 0], [3, 0, 0], [1.4, 0, 0]], np.float32)
    assert g.update(verts, None, [0]) is False
    assert g.update(verts, None, [0, 1, 2, 3, 4]) is True
    assert np.array_equal(g.nodes, np.array([[0, 0, 0], [1, 0, 0], [2, 0, 0], [3, 0, 0]], np.float32))
    # large: 90 000 vertices, graph on one half, ~40 000 candidates farthest first, against the oracle
    rng = np.random.default_rng(7)
    verts, faces = _big_mesh(rng)
    onodes, _ = oracle.sample_nodes(verts[verts[:, 0] < 0.0], np.ones((verts[:, 0] < 0.0).sum(), bool), 0.05, False)
    dist, _ = oracle.nearest_node(verts, onodes)
    cand = np.where(dist > 0.1)[0]
    cand = cand[np.argsort(dist[cand], kind="stable")[::-1]].astype(np.int32)
    acc = oracle.greedy_update(verts, cand, onodes, 0.05)
    g = _graph(onodes.copy(), 0.05)
    assert g.update(verts, None, cand) is True
    assert len(acc) > 100 and np.array_equal(g.nodes[len(onodes):], verts[acc])
    # same candidates in a shuffled order, with duplicates
    cand2 = np.concatenate([rng.permutation(cand)[:5000], cand[:100]]).astype(np.int32)
    acc2 = oracle.greedy_update(verts, cand2, onodes, 0.05)
    g = _graph(onodes.copy(), 0.05)
    g.update(verts, None, cand2)
    assert np.array_equal(g.nodes[len(onodes):], verts[acc2])


def test_find_unreachable_nodes_and_update_graph(gg, oracle):
    """WarpField.find_unreachable_nodes (warpfield.py:455-485) + update_graph (:487-582) through the device operators."""
    from occlusionfusion_b200 import TSDFVolume, WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
    nodes = gg["upd_nodes_before"]
    wf = WarpField(_graph(nodes.copy(), float(gg["coverage"])), vol, None)
    got = np.asarray(wf.find_unreachable_nodes(gg["mesh_verts"]))
    dist, idx = oracle.nearest_node(gg["mesh_verts"], nodes)
    want = np.where(dist > 2 * float(gg["coverage"]))[0]
    want = want[np.argsort(dist[want])[::-1]]
    assert np.array_equal(got, want)
    assert set(got.tolist()) == set(gg["upd_cand"].tolist())            # the reference's own candidate set
    assert len(wf.find_unreachable_nodes(nodes)) == 0                   # every node reaches itself -> [] (warpfield.py:472)
    # the device distances and indices themselves
    import torch
    from occlusionfusion_b200 import _capi
    pts = torch.from_numpy(gg["mesh_verts"]).cuda()
    nd = torch.from_numpy(nodes).cuda()
    d = torch.empty(len(pts), dtype=torch.float32, device="cuda")
    i = torch.empty(len(pts), dtype=torch.int32, device="cuda")
    _capi.check(_capi.load().ofx_nearest_node(_capi.ptr(pts), len(pts), _capi.ptr(nd), len(nd), _capi.ptr(d), _capi.ptr(i),
                                              _capi.stream_ptr()))
    assert np.array_equal(d.cpu().numpy(), dist) and np.array_equal(i.cpu().numpy(), idx)
    assert np.abs(dist - gg["upd_nearest_dist"]).max() <= 1e-7           # vs the reference run (float64 tree cast to f32)
    # update_graph end to end (erode -> unreachable among the non-eroded -> greedy insert -> re-skin), expected result
    # assembled from the oracle's stages with the reference's indexing (warpfield.py:500-516)
    verts, faces, c = gg["mesh_verts"], gg["mesh_faces"], float(gg["coverage"])
    ne = oracle.erode_mesh(len(verts), faces, 1, 3)
    dsub, _ = oracle.nearest_node(verts[ne], nodes)
    cand = np.where(dsub > 2 * c)[0]
    cand = cand[np.argsort(dsub[cand])[::-1]]
    acc = oracle.greedy_update(verts, cand, nodes, c)
    wf.frame_id = vol.frame_id = 0
    assert wf.update_graph((verts, faces)) is True
    assert len(acc) > 20 and np.array_equal(wf.graph.nodes, np.concatenate([nodes, verts[acc]]))
    assert wf.rotations.shape == (len(wf.graph.nodes), 3, 3) and wf.translations.shape == wf.graph.nodes.shape
    a, wt, v = wf.skin_tsdf()                                            # the volume was re-skinned against the new table
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    ov.set_graph(wf.graph.nodes, c, wf.graph_neighbours)
    oa, owt, ovl = ov.skin_tsdf()
    assert np.array_equal(a, oa) and np.array_equal(wt, owt) and np.array_equal(v, ovl)


def test_truncated_region_golden_and_oracle(gg, oracle):
    import torch
    from occlusionfusion_b200 import TSDFVolume
    k4 = load_golden("ref_k4")
    for key in ("tsdf_0", "tsdf_2"):                                     # Dz = 36: the float4 path
        for md in (1.2, 0.5):
            got = TSDFVolume.compute_truncated_region(k4[key], md)
            want = np.unpackbits(gg[f"trunc_{key}_{md}"])[:got.size].astype(bool).reshape(got.shape)
            assert got.dtype == bool and np.array_equal(got, want), (key, md)
    t = gg["trunc_smooth_in"]
    for md in (1.2, 0.3):
        got = TSDFVolume.compute_truncated_region(t, md)
        want = np.unpackbits(gg[f"trunc_smooth_{md}"])[:got.size].astype(bool).reshape(got.shape)
        assert np.array_equal(got, want), md
    # ragged Dz (scalar path), degenerate dims, and a CUDA tensor in -> CUDA tensor out
    rng = np.random.default_rng(8)
    for dims in ((17, 19, 23), (3, 3, 3), (2, 5, 8), (1, 1, 1), (40, 36, 64)):
        x = np.clip(rng.normal(0, 0.6, dims), -1, 1).astype(np.float32)
        got = TSDFVolume.compute_truncated_region(x, 0.8)
        assert np.array_equal(got, oracle.truncated_region(x, 0.8)), dims
    xt = torch.from_numpy(x).cuda()
    out = TSDFVolume.compute_truncated_region(xt, 0.8)
    assert out.is_cuda and out.dtype == torch.bool and np.array_equal(out.cpu().numpy(), oracle.truncated_region(x, 0.8))


def test_truncated_region_of_integrated_volume(oracle):
    """After real integrations: the mask of the device volume equals the oracle's mask of the oracle's volume."""
    from occlusionfusion_b200 import TSDFVolume
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    vol.integrate({"im": w.frame(0), "id": 0})
    ov.integrate(w.frame(0), True)
    got = vol.get_truncated_region().cpu().numpy()
    want = oracle.truncated_region(ov.tsdf, 1.2)
    assert want.sum() > 100 and np.array_equal(got, want)


def _hemisphere_mesh(w, nt=40, nphi=80):
    """The camera-facing hemisphere of the band scene as a (theta, phi) grid mesh."""
    th = np.linspace(0.02, np.pi / 2 - 0.02, nt)
    ph = np.linspace(0, 2 * np.pi, nphi, endpoint=False)
    T, P = np.meshgrid(th, ph, indexing="ij")
    d = np.stack([np.sin(T) * np.cos(P), np.sin(T) * np.sin(P), -np.cos(T)], -1).reshape(-1, 3)
    verts = (w.centre + w.radius * d).astype(np.float32)
    i, j = np.meshgrid(np.arange(nt - 1), np.arange(nphi), indexing="ij")
    a, b = (i * nphi + j).ravel(), (i * nphi + (j + 1) % nphi).ravel()
    faces = np.concatenate([np.stack([a, b, a + nphi], 1), np.stack([b, b + nphi, a + nphi], 1)]).astype(np.int32)
    return verts, faces


def test_sequence_with_update_graph_on_device(oracle):
    """BASELINE config 3 in small, with the node insertion itself on the device: the graph starts on 60 % of the scene,
    WarpField.update_graph (erode -> unreachable vertices -> greedy insertion -> re-skin) extends it from the canonical mesh
    between frames, and every frame stays bit-identical to the oracle driven through the same stages."""
    from occlusionfusion_b200 import TSDFVolume, WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    verts, faces = _hemisphere_mesh(w)
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    nodes = w.nodes[: int(0.6 * w.M)].copy()
    wf = WarpField(_graph(nodes, w.coverage), vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf
    ov.set_graph(nodes, w.coverage, w.K)
    grown = []
    for f in range(7):
        if f in (3, 5):                                          # between frames: warp field and volume agree on the frame id
            ne = oracle.erode_mesh(len(verts), faces, 1, 3)
            d, _ = oracle.nearest_node(verts[ne], wf.graph.nodes)
            cand = np.where(d > 2 * w.coverage)[0]
            cand = cand[np.argsort(d[cand])[::-1]]
            acc = oracle.greedy_update(verts, cand, wf.graph.nodes, w.coverage)
            want = np.concatenate([wf.graph.nodes, verts[acc]]) if len(acc) else wf.graph.nodes
            assert wf.update_graph((verts, faces)) is (len(acc) > 0)
            assert np.array_equal(wf.graph.nodes, want)
            grown.append(len(acc))
            ov.set_graph(wf.graph.nodes, w.coverage, w.K)
        im = w.frame(f)
        if f > 0:
            est = w.estimate(f, wf.graph.nodes)
            wf.update_transformations(est)
            ov.set_transforms(est["node_rotations"], wf.translations)
        vol.integrate({"im": im, "id": f})
        n = ov.integrate(im, f == 0)
        assert vol.last_stats()["updated"] == n, f
    assert grown[0] > 5 and grown[1] == 0                        # the first update covers the scene, the second finds nothing
    t, c, wt = vol.get_volume()
    assert np.array_equal(wt, ov.weight) and np.array_equal(c, ov.colour) and np.array_equal(t, ov.tsdf)


def test_graph_capi_edge_cases():
    """Empty inputs and bad arguments at the C ABI: nothing launches, counts come back zero, errors carry a message."""
    import ctypes
    import torch
    from occlusionfusion_b200 import _capi
    lib, S, P = _capi.load(), _capi.stream_ptr, _capi.ptr
    dev = torch.device("cuda")
    pts = torch.rand((8, 3), device=dev)
    n_out = torch.full((1,), 7, dtype=torch.int32, device=dev)
    acc = torch.empty(8, dtype=torch.int32, device=dev)
    assert lib.ofx_nearest_node(None, 0, P(pts), 8, None, None, S()) == 0                       # no points
    assert lib.ofx_nearest_node(P(pts), 8, P(pts), 0, P(pts), None, S()) != 0                   # no nodes
    assert b"ofx_nearest_node" in lib.ofx_last_error()
    assert lib.ofx_greedy_insert(P(pts), 8, None, 0, P(pts), 8, 0.05, None, P(n_out), S()) == 0  # no candidates
    assert int(n_out.item()) == 0
    n_out.fill_(7)
    assert lib.ofx_sample_nodes(None, 0, None, ctypes.c_float(0.05), 1, None, None, P(n_out), S()) == 0
    assert int(n_out.item()) == 0
    assert lib.ofx_sample_nodes(P(pts), 8, None, ctypes.c_float(0.05), 1, P(acc), None, P(n_out), S()) != 0   # mask required
    ne = torch.full((8,), 9, dtype=torch.uint8, device=dev)
    assert lib.ofx_erode_mesh(None, 0, 8, 1, 3, P(ne), S()) == 0 and int(ne.sum()) == 0           # no faces: everything eroded
    bad = torch.tensor([[0, 1, 99], [0, 1, 2], [1, 2, 3], [0, 2, 3], [0, 1, 3]], dtype=torch.int32, device=dev)
    assert lib.ofx_erode_mesh(P(bad), 5, 8, 0, 3, P(ne), S()) == 0                                # out-of-range face ignored
    assert ne.cpu().tolist() == [1, 1, 1, 1, 0, 0, 0, 0]
    assert lib.ofx_truncated_region(None, 0, 4, 4, 1.2, None, S()) == 0
    assert lib.ofx_truncated_region(None, 4, 4, 4, 1.2, None, S()) != 0
    # no existing nodes (M = 0): plain greedy coverage sampling in candidate order
    v = torch.tensor([[0, 0, 0], [0.01, 0, 0], [1, 0, 0], [1, 0.02, 0]], dtype=torch.float32, device=dev)
    cand = torch.tensor([3, 0, 1, 2], dtype=torch.int32, device=dev)
    assert lib.ofx_greedy_insert(P(v), 4, P(cand), 4, None, 0, 0.05, P(acc), P(n_out), S()) == 0
    assert int(n_out.item()) == 2 and acc[:2].cpu().tolist() == [3, 0]


def test_create_graph_from_mesh_end_to_end(gg, oracle):
    """EDGraph.create_graph_from_mesh (embedded_deformation_graph.py:174-256): device erosion + node sampling, then the host
    half (geodesic edges, clean-up, clusters, pyramid), against the same pipeline assembled from the separately pinned parts."""
    from occlusionfusion_b200.graph_proc import compute_clusters_c, compute_edges_geodesic_c, node_and_edge_clean_up_c
    verts, faces = gg["mesh_verts"], gg["mesh_faces"]
    g = _graph(np.zeros((0, 3), np.float32), float(gg["coverage"]))
    g.create_graph_from_mesh(verts, faces)
    idx0 = gg["sample_idx_1"]                                            # the reference's own sample_nodes on this mesh
    e = -np.ones((len(idx0), 8), np.int32)
    w = np.zeros((len(idx0), 8), np.float32)
    d = np.zeros((len(idx0), 8), np.float32)
    compute_edges_geodesic_c(verts, np.ones(len(verts), bool), faces, idx0, 8, float(gg["coverage"]), e, w, d, None, True, True)
    mask = np.ones((len(idx0), 1), bool)
    node_and_edge_clean_up_c(e, mask)
    keep = mask.reshape(-1)
    assert np.array_equal(g.node_indices.reshape(-1), idx0[keep]) and np.array_equal(g.nodes, gg["sample_pos_1"][keep])
    assert g.edges.shape == (keep.sum(), 8) and (g.edges[:, :2] >= 0).all() and g.edges.max() < keep.sum()
    if keep.all():
        assert np.array_equal(g.edges, e) and np.array_equal(g.edges_weights, w) and np.array_equal(g.edges_distances, d)
    cl = -np.ones((len(g.edges), 1), np.int32)
    sizes = compute_clusters_c(g.edges, cl)
    assert np.array_equal(g.clusters, cl) and sum(sizes) == len(g.nodes)
    pyd = g.get_graph_pyramid()
    assert pyd["nn_index_l0"] is g.edges and pyd["nn_index_l1"].shape[1] == 6 and pyd["nn_index_l3"].shape[1] == 3
    assert len(pyd["up_sample_idx1"]) == len(g.nodes) and len(pyd["down_sample_idx1"]) == len(pyd["nn_index_l1"])


def test_full_update_matches_reference(gg):
    """The whole of the reference's EDGraph.update (embedded_deformation_graph.py:496-605) run here with its own Python and
    C++: insertion, nearest vertices, geodesic edges, clean-up, clusters, pyramid."""
    g = _graph(gg["upd_nodes_before"].copy(), float(gg["coverage"]))
    assert g.update(gg["mesh_verts"], gg["mesh_faces"], gg["upd_cand"]) is True
    assert np.array_equal(g.nodes, gg["full_nodes"])
    assert np.array_equal(np.asarray(g.node_indices).reshape(-1), gg["full_node_indices"].reshape(-1))
    assert np.array_equal(g.edges, gg["full_edges"]) and np.array_equal(g.edges_distances, gg["full_dist"])
    assert np.array_equal(g.edges_weights, gg["full_weights"]) and np.array_equal(g.clusters, gg["full_clusters"])
    keys = [k[8:] for k in gg.files if k.startswith("fullpyd_")]
    assert sorted(g.pyd) == sorted(keys)
    for k in keys:
        assert np.array_equal(np.asarray(g.pyd[k]), gg["fullpyd_" + k]), k


def test_pixel_anchors_euclidean_match_reference(gg):
    """compute_pixel_anchors_euclidean (csrc/cpu/graph_proc.cpp:610-709) on the device against the reference's C++: anchors
    identical (duplicated nodes exercise the later-node-first tie rule), weights within 1e-6 (std::exp on a float there,
    float64 exp rounded once here)."""
    from occlusionfusion_b200.graph_proc import compute_pixel_anchors_euclidean_c
    for name, nodes in (("all", gg["pix_nodes"]), ("three", gg["pix_nodes"][:3])):
        a, w = np.zeros(0, np.int32), np.zeros(0, np.float32)
        compute_pixel_anchors_euclidean_c(nodes, gg["pix_image"], float(gg["coverage"]), a, w)
        assert a.shape == gg[f"pix_{name}_anchors"].shape and np.array_equal(a, gg[f"pix_{name}_anchors"])
        assert np.abs(w - gg[f"pix_{name}_weights"]).max() <= 1e-6
        assert (w == gg[f"pix_{name}_weights"]).mean() > 0.99
    assert (gg["pix_all_anchors"] >= 120).sum() > 50 and (gg["pix_all_anchors"][..., 0] == -1).mean() > 0.05
    assert (gg["pix_three_anchors"][..., 3] == -1).all()                 # three nodes: the fourth slot stays empty
