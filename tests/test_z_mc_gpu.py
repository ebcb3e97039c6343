"""GPU: marching cubes on the device (ofx_mc_*; TSDFVolume.marching_cubes / get_mesh) against the CPU oracle's restatement
of the same contract (oracle/ofx_oracle_mc.c): vertices, faces and colours identical.  Parity with the reference itself is
UNPINNED for this step: the reference calls scikit-image (Lewiner's variant), which is not installed here."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _np_normals(vol, keys, level=0.0):
    W, H, D = vol.shape
    lower, axis = keys // 3, keys % 3
    x, y, z = lower // (H * D), (lower // D) % H, lower % D

    def grad(x, y, z):
        xp, xm = np.minimum(x + 1, W - 1), np.maximum(x - 1, 0)
        yp, ym = np.minimum(y + 1, H - 1), np.maximum(y - 1, 0)
        zp, zm = np.minimum(z + 1, D - 1), np.maximum(z - 1, 0)
        return np.stack([vol[xp, y, z] - vol[xm, y, z], vol[x, yp, z] - vol[x, ym, z], vol[x, y, zp] - vol[x, y, zm]], 1)

    g0 = grad(x, y, z)
    g1 = grad(x + (axis == 0), y + (axis == 1), z + (axis == 2))
    stride = np.array([H * D, D, 1])[axis]
    v0, v1 = vol.reshape(-1)[lower], vol.reshape(-1)[lower + stride]
    t = ((np.float32(level) - v0) / (v1 - v0)).astype(np.float32)[:, None]
    g = (g0 + t * (g1 - g0)).astype(np.float32)
    s = (g[:, 0] * g[:, 0] + g[:, 1] * g[:, 1]) + g[:, 2] * g[:, 2]
    return g * np.where(s > 0, np.float32(1) / np.sqrt(s), 0).astype(np.float32)[:, None]


def _check(vol, mask, oracle, colour=None):
    from occlusionfusion_b200 import TSDFVolume
    v, f, n, rgb = TSDFVolume.marching_cubes(vol, mask, 0.0, colour)
    ov, of, okeys = oracle.marching_cubes(vol, mask, 0.0)
    assert np.array_equal(f.cpu().numpy(), of) and np.array_equal(v.cpu().numpy(), ov)
    if len(ov):
        assert np.abs(n.cpu().numpy() - _np_normals(vol, okeys)).max() < 1e-6
    if colour is not None and len(ov):
        ind = np.round(ov).astype(int)                                   # tsdf.py:797-807
        c = colour[ind[:, 0], ind[:, 1], ind[:, 2]]
        b = np.floor(c / 65536)
        g = np.floor((c - b * 65536) / 256)
        r = c - b * 65536 - g * 256
        assert np.array_equal(rgb.cpu().numpy(), np.floor(np.asarray([r, g, b])).T.astype(np.uint8))
    return ov, of


def test_sphere_is_watertight_and_matches_oracle(oracle):
    n = 48
    x, y, z = np.meshgrid(*(np.arange(n),) * 3, indexing="ij")
    centre = np.array([22.3, 24.1, 21.7])
    vol = (np.sqrt((x - centre[0]) ** 2 + (y - centre[1]) ** 2 + (z - centre[2]) ** 2) - 14.4).astype(np.float32)
    v, f = _check(vol, None, oracle)
    e = np.sort(np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]), axis=1)
    _, counts = np.unique(e, axis=0, return_counts=True)
    assert (counts == 2).all() and len(v) - len(counts) + len(f) == 2          # closed manifold, Euler characteristic 2
    assert np.abs(np.linalg.norm(v - centre, axis=1) - 14.4).max() < 0.02
    a, b, c = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    assert (((np.cross(b - a, c - a)) * ((a + b + c) / 3 - centre)).sum(1) > 0).all()   # normals from inside to outside


def test_random_fields_masks_and_ragged_dims(oracle):
    rng = np.random.default_rng(3)
    for dims in ((14, 13, 15), (9, 33, 70), (2, 2, 2), (1, 5, 5), (31, 8, 260)):
        vol = rng.normal(size=dims).astype(np.float32)
        colour = np.floor(rng.random(dims) * 16777215).astype(np.float32)
        _check(vol, None, oracle, colour)
        _check(vol, rng.random(dims) < 0.4, oracle, colour)
    _check(np.ones((8, 8, 8), np.float32), None, oracle)                       # no surface at all


def test_get_mesh_of_integrated_volume(oracle):
    """TSDFVolume.get_mesh (tsdf.py:779-809): truncated region -> marching cubes -> world coordinates + colours."""
    from occlusionfusion_b200 import TSDFVolume
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    vol.integrate({"im": w.frame(0), "id": 0})
    ov.integrate(w.frame(0), True)
    verts, faces, norms, colors = vol.get_canonical_model()
    region = oracle.truncated_region(ov.tsdf, 1.2)
    overts, ofaces, _ = oracle.marching_cubes(ov.tsdf, region, 0.0)
    assert len(ofaces) > 200 and np.array_equal(faces, ofaces) and faces.dtype == np.int32
    assert np.array_equal(verts, overts * np.float32(w.voxel_size) + w.origin)
    assert colors.dtype == np.uint8 and colors.shape == verts.shape and norms.shape == verts.shape
    # the surface lies on the rendered sphere (band scene): projective TSDF: within a voxel or two of its radius, worst at the silhouette
    r = np.linalg.norm(verts - w.centre, axis=1)
    assert np.abs(r - w.radius).max() < 2.5 * w.voxel_size and np.median(np.abs(r - w.radius)) < 0.5 * w.voxel_size
    assert vol.get_canonical_model() is vol.canonical_model                     # cached (tsdf.py:813)


def test_update_graph_standalone_from_the_volume(oracle):
    """WarpField.update_graph() with no arguments (warpfield.py:487-582): the canonical mesh comes from the volume itself
    (truncated region -> marching cubes), nodes are inserted where the graph does not reach, the volume is re-skinned, and the
    following frames stay bit-identical to the oracle driven through the same stages."""
    from occlusionfusion_b200 import EDGraph, TSDFVolume, WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    nodes = w.nodes[: int(0.6 * w.M)].copy()
    wf = WarpField(EDGraph(nodes=nodes, graph_generation_parameters={"node_coverage": w.coverage}), vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf
    ov.set_graph(nodes, w.coverage, w.K)
    vol.integrate({"im": w.frame(0), "id": 0})
    ov.integrate(w.frame(0), True)
    # expected insertion, stage by stage on the oracle's volume
    overts, ofaces, _ = oracle.marching_cubes(ov.tsdf, oracle.truncated_region(ov.tsdf, 1.2), 0.0)
    verts = overts * np.float32(w.voxel_size) + w.origin
    ne = oracle.erode_mesh(len(verts), ofaces, 1, 3)
    d, _ = oracle.nearest_node(verts[ne], nodes)
    cand = np.where(d > 2 * w.coverage)[0]
    cand = cand[np.argsort(d[cand])[::-1]]
    acc = oracle.greedy_update(verts, cand, nodes, w.coverage)
    assert len(acc) > 3
    assert wf.update_graph() is True
    new_nodes = wf.graph.nodes
    inserted = {tuple(p) for p in verts[acc].tolist()}
    assert inserted.issuperset({tuple(p) for p in new_nodes.tolist()} - {tuple(p) for p in nodes.tolist()})
    if len(new_nodes) == len(nodes) + len(acc):                     # no node lost in the edge rebuild: exactly the greedy result
        assert np.array_equal(new_nodes, np.concatenate([nodes, verts[acc]]))
    assert wf.rotations.shape == (len(new_nodes), 3, 3) and hasattr(wf.graph, "edges") and wf.graph.edges.shape[0] == len(new_nodes)
    ov.set_graph(new_nodes, w.coverage, w.K)
    for f in (1, 2):
        est = w.estimate(f, new_nodes)
        wf.update_transformations(est)
        ov.set_transforms(est["node_rotations"], wf.translations)
        vol.integrate({"im": w.frame(f), "id": f})
        assert vol.last_stats()["updated"] == ov.integrate(w.frame(f), False)
    t, c, wt = vol.get_volume()
    assert np.array_equal(wt, ov.weight) and np.array_equal(c, ov.colour) and np.array_equal(t, ov.tsdf)


def test_deformed_model_and_clear(oracle):
    """get_deformed_model (tsdf.py:830-849) = canonical mesh through WarpField.deform_mesh; clear() drops the caches."""
    from occlusionfusion_b200 import TSDFVolume, WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
    wf = WarpField(w.graph(), vol, None)
    vol.warpfield = wf
    vol.integrate({"im": w.frame(0), "id": 0})
    assert vol.get_deformed_model() is vol.get_canonical_model()                # source frame: the canonical model itself
    vol.clear()
    assert not hasattr(vol, "canonical_model") and not hasattr(vol, "deformed_model")
    est = w.estimate(1)
    wf.update_transformations(est)
    vol.integrate({"im": w.frame(1), "id": 1})
    dv, df, dn, dc = vol.get_deformed_model()
    cv, cf, cn, cc = vol.get_canonical_model()
    a, wt, valid = oracle.skin(cv, w.nodes, w.K, w.coverage)
    assert np.array_equal(dv, oracle.warp(cv, a, wt, valid, est["node_rotations"], wf.translations))
    assert df is cf and dc is cc and dn.shape == cn.shape and valid.mean() > 0.9


def test_fusion_py_call_sequence_against_the_drop_ins(oracle, tmp_path):
    """The calls DynamicFusion makes on the three classes, in its order and with its argument order
    (fusion_with_occlusion/fusion.py:57-83 create_tsdf, :176-179 register_frame): TSDFVolume(bbox, max_depth, intrinsics,
    opt, vis) -> integrate(source_data) -> EDGraph(tsdf, vis, source_data) -> WarpField(graph, tsdf, vis) ->
    update_transformations(estimated) -> integrate(target_frame_data).  The graph comes from the volume's own mesh and
    <datadir>/graph_config.json; the warped frames stay bit-identical to the oracle given the same node table."""
    import json
    import os
    import types
    from occlusionfusion_b200 import EDGraph, TSDFVolume, WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny", dim=[64, 64, 96], voxel=0.0125)
    opt = types.SimpleNamespace(voxel_dim=w.fopt.voxel_dim, source_frame=0, skip_rate=1, datadir=str(tmp_path), gpu=True)
    vis = None
    source_data = {"im": w.frame(0), "id": 0, "intrinsics": w.intr}
    tsdf = TSDFVolume(w.bbox, w.max_depth, source_data["intrinsics"], opt, vis)          # fusion.py:57
    tsdf.integrate(source_data)                                                           # fusion.py:71
    graph = EDGraph(tsdf, vis, source_data)                                               # fusion.py:74
    tsdf.graph = graph                                                                    # fusion.py:76
    warpfield = WarpField(graph, tsdf, vis)                                               # fusion.py:83
    tsdf.warpfield = warpfield                                                            # fusion.py:86
    # what the constructor left behind (embedded_deformation_graph.py:34-52)
    cfg = json.load(open(os.path.join(str(tmp_path), "graph_config.json")))
    assert cfg["node_coverage"] == 0.05 and cfg["num_neighbours"] == 8 and graph.graph_generation_parameters == cfg
    assert os.path.isdir(os.path.join(str(tmp_path), "results", "updated_graph"))
    assert graph.tsdf is tsdf and tsdf.graph is graph
    M = graph.nodes.shape[0]
    assert M > 20 and graph.nodes.dtype == np.float32 and graph.edges.shape == (M, 8) and "nn_index_l3" in graph.pyd
    # nodes lie on the surface mesh, node_coverage apart
    verts = tsdf.get_canonical_model()[0]
    assert np.array_equal(graph.nodes, verts[graph.node_indices.reshape(-1)])
    d = np.linalg.norm(graph.nodes[:, None] - graph.nodes[None], axis=2) + 10 * np.eye(M)
    assert d.min() > 0.05
    # a graph_config.json that is already there is read, not overwritten
    cfg["node_coverage"] = 0.08
    json.dump(cfg, open(os.path.join(str(tmp_path), "graph_config.json"), "w"))
    assert EDGraph(tsdf, vis, source_data).nodes.shape[0] < M
    tsdf.graph = graph
    # registered frames (fusion.py:176-179)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    ov.integrate(w.frame(0), True)
    ov.set_graph(graph.nodes, warpfield.node_coverage, warpfield.graph_neighbours)
    for f in (1, 2):
        estimated_transformations = w.estimate(f, graph.nodes)
        warpfield.update_transformations(estimated_transformations)                       # fusion.py:176
        target_frame_data = {"im": w.frame(f), "id": f, "intrinsics": w.intr}
        tsdf.integrate(target_frame_data)                                                 # fusion.py:179
        ov.set_transforms(estimated_transformations["node_rotations"], warpfield.translations)
        assert tsdf.last_stats()["updated"] == ov.integrate(w.frame(f), False) > 0
    t, c, wt = tsdf.get_volume()
    assert np.array_equal(wt, ov.weight) and np.array_equal(c, ov.colour) and np.array_equal(t, ov.tsdf)
