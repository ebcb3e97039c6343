"""GPU parity: the CUDA path (through the drop-in classes, i.e. through the C ABI of libofx.so)
against (1) the golden vectors produced by the reference's own Python and (2) the C oracle on
seeded synthetic inputs.  Bars (BASELINE.json north_star): anchors and integer weight counts
bit-exact; TSDF, skinning weights and colours within 1e-5 absolute.  Against the oracle the CUDA
path is expected to be bit-identical throughout (same op order), which the tests assert."""
import types

import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _mk_volume(g_or_w, dims=None, **kw):
    from occlusionfusion_b200 import TSDFVolume
    if isinstance(g_or_w, dict) or hasattr(g_or_w, "files"):
        g = g_or_w
        fopt = types.SimpleNamespace(voxel_dim=[int(d) for d in g["voxel_dim"]], source_frame=0, skip_rate=1)
        return TSDFVolume(tuple(g["bbox"]), float(g["max_depth"]), g["intr"], fopt, None, **kw)
    w = g_or_w
    return TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None, **kw)


def _graph(nodes, coverage):
    from occlusionfusion_b200 import EDGraph
    return EDGraph(nodes=nodes, graph_generation_parameters={"node_coverage": float(coverage)})


@pytest.mark.parametrize("name", ["ref_k4", "ref_k8", "ref_m3"])
def test_golden_sequence(name, oracle):
    """The reference's own run, frame by frame: geometry, frame packing, warp, volumes."""
    from occlusionfusion_b200 import WarpField
    g = load_golden(name)
    vol = _mk_volume(g)
    assert np.array_equal(vol._vol_dim, g["voxel_dim"])
    assert vol._voxel_size == float(g["voxel_size"])
    assert np.array_equal(vol._vol_origin, g["vol_origin"])
    assert np.array_equal(vol.world_pts, g["world_pts"])

    K = min(int(g["K"]), len(g["nodes"]))
    ov = oracle.OracleVolume(g["voxel_dim"], g["vol_origin"], g["voxel_size"], g["intr"], g["trunc"])
    ov.set_graph(g["nodes"], float(g["coverage"]), K)

    vol.integrate({"im": g["frames"][0], "id": 0})
    ov.integrate(g["frames"][0], True)
    assert np.array_equal(vol.color_im, g["color_im_0"]) and np.array_equal(vol.depth_im, g["depth_im_0"])
    t, c, w = vol.get_volume()
    assert np.array_equal(w, g["weight_0"]) and np.array_equal(c, g["color_0"])
    assert np.array_equal(t, g["tsdf_0"])          # identity warp: no float32 warp in the way -> exact
    assert vol.last_stats()["updated"] == int((g["weight_0"] != 0).sum())

    wf = WarpField(_graph(g["nodes"], g["coverage"]), vol, None)
    wf.graph_neighbours = K
    vol.warpfield = wf
    for f in range(1, len(g["frames"])):
        est = dict(node_rotations=g[f"R_{f}"], node_translations=g[f"t_{f}"],
                   deformed_nodes_to_target=g["nodes"] + g[f"t_{f}"], target_frame_id=f)
        wf.update_transformations(est)
        assert np.array_equal(wf.translations, g[f"T_{f}"])
        vol.integrate({"im": g["frames"][f], "id": f})
        ov.set_transforms(g[f"R_{f}"], g[f"T_{f}"])
        n_upd = ov.integrate(g["frames"][f], False)
        t, c, w = vol.get_volume()
        # against the reference's run
        assert np.array_equal(w, g[f"weight_{f}"]), "integer weight counts must be bit-exact"
        assert np.array_equal(c, g[f"color_{f}"])
        assert np.abs(t - g[f"tsdf_{f}"]).max() <= 1e-5
        # against the oracle: bit-identical
        assert np.array_equal(t, ov.tsdf) and np.array_equal(w, ov.weight) and np.array_equal(c, ov.colour)
        st = vol.last_stats()
        assert st["updated"] == n_upd
        assert st["weight_sum"] == int(ov.weight[ov.weight != (g[f"weight_{f-1}"])].sum())
        warped, valid = wf.deform_tsdf()
        assert np.array_equal(valid, g[f"valid_{f}"])
        assert np.abs(warped - g[f"warped_{f}"]).max() <= 1e-6
        ow, _ = ov.deform_tsdf()
        assert np.array_equal(warped, ow)
    a, wts, v = wf.skin_tsdf()
    oa, owt, ovl = ov.skin_tsdf()
    assert np.array_equal(a, oa) and np.array_equal(wts, owt) and np.array_equal(v, ovl)
    assert np.array_equal(v, g["skin_valid"])
    # weights against the reference's run, on the rows whose anchor ORDER agrees (the golden run's KD-tree is scipy's
    # float64 one; pykdtree's float32 tree is absent here, so the order of two anchors at float32-equal distances is
    # not observable: the K = 8 case has two such rows, same anchor set, two entries swapped)
    same = (a == g["anchors"]).all(axis=1)
    assert np.array_equal(np.sort(a, axis=1), np.sort(g["anchors"], axis=1)), "anchor SETS must agree on every voxel"
    assert (~same).sum() <= 2 and np.abs(wts[same] - g["weights"][same]).max() <= 1e-6
    assert np.abs(np.sort(wts[~same], axis=1) - np.sort(g["weights"][~same], axis=1)).max(initial=0) <= 1e-6


@pytest.mark.parametrize("name", ["ref_k4", "ref_k8", "ref_m3"])
def test_point_operators_golden(name, oracle):
    """skin / deform / check_visibility on free point sets (mesh vertices, graph nodes)."""
    from occlusionfusion_b200 import WarpField
    g = load_golden(name)
    vol = _mk_volume(g)
    last = len(g["frames"]) - 1
    vol.update(g["frames"][last], 0)
    wf = WarpField(_graph(g["nodes"], g["coverage"]), vol, None)
    K = min(int(g["K"]), len(g["nodes"]))
    wf.graph_neighbours = K
    sub = g["nodes"][: max(int(g["K"]), len(g["nodes"]) // 2)]
    a, w, v = wf.skin(g["skin_q"], nodes=sub)
    assert a.dtype == np.int32 and w.dtype == np.float32 and v.dtype == bool
    assert np.array_equal(a, g["skin_q_anchors"]) and np.array_equal(v, g["skin_q_valid"])
    assert np.abs(w - g["skin_q_weights"]).max() <= 1e-6
    oa, ow, ovl = oracle.skin(g["skin_q"], sub, K, float(g["coverage"]))
    assert np.array_equal(a, oa) and np.array_equal(w, ow) and np.array_equal(v, ovl)
    vis, diff = vol.check_visibility(g["skin_q"])
    assert np.array_equal(vis, g["vis_valid"]) and np.array_equal(diff, g["vis_diff"])
    # deform(): anchors index the sub-table
    R, T = g[f"R_{last}"][: len(sub)], g[f"T_{last}"][: len(sub)]
    wf.rotations, wf.translations = R, T
    out = wf.deform(g["skin_q"], a, w, None, v)
    assert np.array_equal(out, oracle.warp(g["skin_q"], a, w, v, R, T))
    assert np.array_equal(out[~v], g["skin_q"][~v])               # invalid points pass through


def test_reference_side_files_normals_and_volume_pickle(oracle, tmp_path):
    """tests/golden/ref_misc.npz (oracle/make_golden_misc.py: the reference's own Python run here): the .npy side files of
    update_transformations / get_visible_nodes byte for byte, get_visible_nodes, the transformation helpers,
    deform_normals, and the volume pickle in both directions (tsdf.py:682-702: (3,Dx,Dy,Dz) = [tsdf, colour, weight])."""
    import pickle
    from occlusionfusion_b200 import TSDFVolume, WarpField
    g = load_golden("ref_misc")
    fopt = types.SimpleNamespace(voxel_dim=[int(d) for d in g["voxel_dim"]], source_frame=0, skip_rate=1, datadir=str(tmp_path))
    vol = TSDFVolume(tuple(g["bbox"]), float(g["max_depth"]), g["intr"], fopt, None)
    vol.integrate({"im": g["frames"][0], "id": 0})
    wf = WarpField(_graph(g["nodes"], g["coverage"]), vol, None)
    wf.graph_neighbours = int(g["K"])
    vol.warpfield = wf
    est = dict(node_rotations=g["R"], node_translations=g["t"], deformed_nodes_to_target=g["nodes"] + g["t"], target_frame_id=1)
    wf.update_transformations(est)
    assert np.array_equal(wf.translations, g["T"])
    got = open(tmp_path / "results" / "deformed_nodes" / "1.npy", "rb").read()
    assert got == g["deformed_nodes_npy"].tobytes(), "deformed_nodes/<id>.npy differs from the file the reference writes"
    Rn, tn = wf.get_transformation_wrt_graph_node()
    assert np.array_equal(Rn, g["R_wrt_node"]) and np.array_equal(tn, g["t_wrt_node"])
    Ro, To = wf.get_transformation_wrt_origin(Rn, tn)
    assert np.array_equal(Ro, g["R_wrt_origin"]) and np.array_equal(To, g["T_wrt_origin"])
    assert np.abs(To - wf.translations).max() <= 1e-6                    # the round trip
    vol.integrate({"im": g["frames"][1], "id": 1})
    vis = vol.get_visible_nodes()
    assert vis.dtype == bool and np.array_equal(vis, g["visible_nodes"])
    assert np.array_equal(vis, oracle.check_visibility(wf.get_deformed_nodes(), vol.depth_im, g["intr"], 0.04)[0])
    got = open(tmp_path / "results" / "visible_nodes" / "1.npy", "rb").read()
    assert got == g["visible_nodes_npy"].tobytes(), "visible_nodes/<id>.npy differs from the file the reference writes"
    # deform_normals: rotate with translation 0, renormalise (warpfield.py:312-345)
    n = len(g["dn_points"])
    dn = wf.deform_normals(g["dn_normals"], g["dn_anchors"], g["dn_weights"], [n, 1, 1], np.ones(n, bool))
    assert dn.shape == (n, 3) and np.abs(dn - g["dn_out"]).max() <= 1e-6
    raw = oracle.warp(g["dn_normals"], g["dn_anchors"], g["dn_weights"], np.ones(n, bool), g["R"], np.zeros_like(g["T"], dtype=np.float32))
    assert np.array_equal(dn, raw / np.linalg.norm(raw, axis=1, keepdims=True))
    assert np.abs(np.linalg.norm(dn, axis=1) - 1).max() <= 1e-6
    # the volume: same values as the reference's run, and the pickle both ways
    t, c, wt = vol.get_volume()
    assert np.array_equal(wt, g["weight"]) and np.array_equal(c, g["colour"]) and np.abs(t - g["tsdf"]).max() <= 1e-5
    mine = tmp_path / "mine.pkl"
    vol.save_volume(str(mine))
    data = pickle.load(open(mine, "rb"))
    assert data.shape == (3,) + tuple(int(d) for d in g["voxel_dim"]) and data.dtype == np.float32
    assert np.array_equal(data[0], t) and np.array_equal(data[1], c) and np.array_equal(data[2], wt)
    theirs = tmp_path / "theirs.pkl"
    theirs.write_bytes(g["volume_pickle"].tobytes())
    ref = pickle.load(open(theirs, "rb"))
    assert ref.shape == data.shape and ref.dtype == data.dtype
    vol2 = TSDFVolume(tuple(g["bbox"]), float(g["max_depth"]), g["intr"], fopt, None)
    vol2.load_volume(str(theirs))
    t2, c2, w2 = vol2.get_volume()
    assert np.array_equal(t2, g["tsdf"]) and np.array_equal(c2, g["colour"]) and np.array_equal(w2, g["weight"])
    vol2.load_volume(str(mine))
    assert all(np.array_equal(x, y) for x, y in zip(vol2.get_volume(), (t, c, wt)))


def _run_pair(w, oracle, n_frames=3, vol_kw=None, obs_weight=1.0):
    """Drives the CUDA classes and the oracle over the same synthetic sequence."""
    from occlusionfusion_b200 import WarpField
    vol = _mk_volume(w, **(vol_kw or {}))
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    ov.set_graph(w.nodes, w.coverage, w.K)
    wf = WarpField(w.graph(), vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf
    stats = []
    for f in range(n_frames):
        im = w.frame(f)
        if f > 0:
            est = w.estimate(f)
            wf.update_transformations(est)
            ov.set_transforms(est["node_rotations"], wf.translations)
        vol.integrate({"im": im, "id": f}, obs_weight)
        n = ov.integrate(im, f == 0, obs_weight)
        stats.append((vol.last_stats(), n))
    return vol, ov, wf, stats


def _assert_bit_equal(vol, ov, sl=slice(None)):
    t, c, w = vol.get_volume()
    assert np.array_equal(w, ov.weight[sl]), "weights differ"
    assert np.array_equal(c, ov.colour[sl]), "colours differ"
    assert np.array_equal(t, ov.tsdf[sl]), f"tsdf differs, max {np.abs(t - ov.tsdf[sl]).max()}"


def test_synthetic_tiny_bit_exact(oracle):
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol, ov, wf, stats = _run_pair(w, oracle, n_frames=4)
    _assert_bit_equal(vol, ov)
    for st, n in stats:
        assert st["updated"] == n
    assert stats[1][1] > 1000, "workload too thin to mean anything"
    a, wt, v = wf.skin_tsdf()
    oa, owt, ovl = ov.skin_tsdf()
    assert np.array_equal(a, oa) and np.array_equal(wt, owt) and np.array_equal(v, ovl)
    assert 0 < stats[1][0]["skin_valid"] <= int(ovl.sum())        # bricks outside the frustum are not visited


def test_synthetic_dense_and_obs_weight(oracle):
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("c1-dense", dim=[48, 40, 64], voxel=0.02, n_nodes=64, H=120, W=160)
    vol, ov, wf, stats = _run_pair(w, oracle, n_frames=3, obs_weight=2.0)
    _assert_bit_equal(vol, ov)
    assert wf.deform_tsdf()[1].all()                                  # every voxel has K anchors
    assert stats[1][1] <= stats[1][0]["skin_valid"] <= w.dims[0] * w.dims[1] * w.dims[2]


def test_obs_weight_not_float32(oracle):
    """obs_weight = 0.3 is not a float32 value: the general float64 weight sum (tsdf.py:373) runs instead
    of the float32 fast path; the source frame takes the generic kernel, warped frames the persistent one."""
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol, ov, wf, stats = _run_pair(w, oracle, n_frames=4, obs_weight=0.3)
    _assert_bit_equal(vol, ov)
    for st, n in stats:
        assert st["updated"] == n


def test_mixed_float32_obs_weights_on_the_packed_path(oracle):
    """Observation weights that are float32 values but not integers (0.625, 1.75) followed by 1.0: the voxel weights
    become non-integers, so w*old is inexact and the colour blend's two roundings (product, then sum) differ from a
    fused multiply-add -- pins the unfused packed product of fuse_pair (ptxas would contract mul.rn.f32x2 +
    add.rn.f32x2).  0.3 in the middle takes the generic kernel (not a float32 value)."""
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("c1-dense", dim=[48, 40, 64], voxel=0.02, n_nodes=64, H=120, W=160)
    vol = _mk_volume(w)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    ov.set_graph(w.nodes, w.coverage, w.K)
    wf = WarpField(w.graph(), vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf
    for f, obs in enumerate([1.0, 0.625, 1.75, 0.3, 1.0, 1.0, 0.625]):
        im = w.frame(f)
        if f > 0:
            est = w.estimate(f)
            wf.update_transformations(est)
            ov.set_transforms(est["node_rotations"], wf.translations)
        vol.integrate({"im": im, "id": f}, obs)
        n = ov.integrate(im, f == 0, obs)
        assert vol.last_stats()["updated"] == n
        _assert_bit_equal(vol, ov)
    t, c, wt = vol.get_volume()
    assert (wt != np.round(wt)).any() and (t < 1.0).any()


def test_medium_dense_certified_paths(oracle):
    """393 k voxels, all skin-valid, three warped frames: large enough that the certified float32 pixel
    (about 1e-4 of the voxels are undecided and take the float64 original), the float32 visibility test
    and the free-space fast path of the running average all see their rare branches; bit-identical."""
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("c1-dense", dim=[64, 64, 96], voxel=0.0125, n_nodes=125, H=240, W=320)
    vol, ov, wf, stats = _run_pair(w, oracle, n_frames=4)
    _assert_bit_equal(vol, ov)
    for st, n in stats:
        assert st["updated"] == n
    assert stats[1][1] > 100000
    t, c, wt = vol.get_volume()
    assert wt.max() == 4.0 and (t < 1.0).any()          # repeated observations and a real surface band


@pytest.mark.parametrize("dims", [[13, 17, 19], [8, 8, 33], [5, 3, 4]])
def test_ragged_dims_scalar_path(dims, oracle):
    """Dz not a multiple of 4 -> unaligned rows -> scalar load/store path; bricks partly empty."""
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny", dim=dims, voxel=0.06, n_nodes=30, H=60, W=80)
    vol, ov, wf, stats = _run_pair(w, oracle, n_frames=3)
    _assert_bit_equal(vol, ov)


@pytest.mark.parametrize("K", [1, 2, 3, 5, 8])
def test_other_k(K, oracle):
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny", K=K, dim=[24, 24, 32], voxel=0.04)
    vol, ov, wf, stats = _run_pair(w, oracle, n_frames=2)
    _assert_bit_equal(vol, ov)
    assert stats[1][1] > 0


def test_sequence_with_node_insertion(oracle):
    """BASELINE config 3 in small: a frame sequence with per-frame graph-node updates and node insertion
    (EDGraph.update appends nodes, warpfield.py:532-533 re-skins every voxel): after each insertion the
    device node table and skin cache are rebuilt and the following frames stay bit-identical to the
    oracle driven with the same node table."""
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol = _mk_volume(w)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    M0 = w.M - 30                                        # start with a subset, insert the rest in three goes
    nodes = w.nodes[:M0].copy()
    wf = WarpField(_graph(nodes, w.coverage), vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf
    ov.set_graph(nodes, w.coverage, w.K)
    n_upd = []
    for f in range(8):
        im = w.frame(f)
        if f > 0:
            if f in (3, 5, 6):                           # node insertion: old nodes keep their indices
                nodes = w.nodes[: len(nodes) + 10].copy()
                wf.graph.nodes = nodes
                wf.rotations = np.tile(np.eye(3, dtype=np.float32), (len(nodes), 1, 1))
                wf.translations = np.zeros((len(nodes), 3), np.float32)
                wf.deformed_nodes = nodes.copy()
                ov.set_graph(nodes, w.coverage, w.K)
            est = w.estimate(f, nodes)
            wf.update_transformations(est)
            ov.set_transforms(est["node_rotations"], wf.translations)
        vol.integrate({"im": im, "id": f})
        n = ov.integrate(im, f == 0)
        assert vol.last_stats()["updated"] == n, f
        n_upd.append(n)
    _assert_bit_equal(vol, ov)
    assert min(n_upd[1:]) > 1000
    a, wt, v = wf.skin_tsdf()
    oa, owt, ovl = ov.skin_tsdf()
    assert np.array_equal(a, oa) and np.array_equal(wt, owt) and np.array_equal(v, ovl)


@pytest.mark.parametrize("x_block", [None, 4])
def test_sharded_equals_whole(x_block, oracle):
    """Two shards (contiguous slabs, and 4-plane blocks dealt round-robin) reproduce the whole volume."""
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    ov = None
    parts = []
    for r in range(2):
        vol, ov, wf, stats = _run_pair(w, oracle, n_frames=3, vol_kw=dict(shard=r, n_shards=2, x_block=x_block))
        parts.append(vol)
    xb = w.dims[0] // 2 if x_block is None else x_block
    for r, vol in enumerate(parts):
        gx = np.array([(lx // xb) * xb * 2 + r * xb + lx % xb for lx in range(w.dims[0] // 2)])
        _assert_bit_equal(vol, ov, gx)


@pytest.mark.parametrize("lookahead", [0, 1, 2])
def test_pipelined_driver_equals_sequential(lookahead, oracle, monkeypatch):
    """sharded.ShardedFusion.pipeline_* (two packets, two streams: the N > 1 bench path) on one shard, without the
    look-ahead, with ofx_prepare_frame (1) and with ofx_stage_frame (2: everything of the next frame but the fused
    kernel staged on the communication stream): same volume and the same per-frame counters as the sequential step(),
    and as the oracle."""
    if lookahead:
        monkeypatch.setenv("OFX_LOOKAHEAD", str(lookahead))
    else:
        monkeypatch.delenv("OFX_LOOKAHEAD", raising=False)
    import torch
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.sharded import CudaShard, ShardedFusion
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    dev = torch.device("cuda", 0)
    vols, counters = [], []
    for mode in ("sequential", "pipelined"):
        vol = _mk_volume(w)
        wf = WarpField(w.graph(), vol, None)
        vol.warpfield = wf
        sf = ShardedFusion(CudaShard(vol), w.dims, 0, 1, w.dims[0], dev)
        sf.set_graph(w.nodes, w.coverage, w.K)
        sf.fill_packet(w.frame(0), np.tile(np.eye(3, dtype=np.float32), (w.M, 1, 1)), np.zeros((w.M, 3), np.float32))
        sf.step(w.H, w.W, identity_warp=True)
        packets = []
        for f in range(1, 5):
            est = w.estimate(f)
            T = -np.einsum("nij,nj->ni", est["node_rotations"], w.nodes) + w.nodes + est["node_translations"]
            sf.fill_packet(w.frame(f), est["node_rotations"], T)
            torch.cuda.synchronize()
            packets.append(sf.buf.clone())
        got = []
        if mode == "sequential":
            for p in packets:
                sf.buf.copy_(p)
                got.append(sf.step(w.H, w.W).clone())
        else:
            sf.pipeline_begin(w.H, w.W, packets[0])
            outs = [sf.pipeline_step(packets[i + 1] if i + 1 < len(packets) else None, last=(i + 1 == len(packets)))
                    for i in range(len(packets))]
            final = sf.pipeline_end()
            torch.cuda.synchronize()
            got = [outs[-1].clone(), final.clone()]           # a step returns the counters of the frame before it
        vols.append(vol.get_volume())
        counters.append([g.cpu().numpy() for g in got])
    for a, b in zip(vols[0], vols[1]):
        assert np.array_equal(a, b)
    # (the look-ahead cull runs before the depth image exists, so it keeps the bricks behind the surface: more bricks
    # and skin-valid voxels VISITED, the same voxels updated)
    cols = [0, 3] if lookahead == 1 else [0, 1, 2, 3]
    assert np.array_equal(counters[0][-1][:, cols], counters[1][-1][:, cols])
    assert np.array_equal(counters[0][-2][:, cols], counters[1][-2][:, cols])
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    ov.set_graph(w.nodes, w.coverage, w.K)
    ov.integrate(w.frame(0), True)
    for f in range(1, 5):
        est = w.estimate(f)
        T = (-np.einsum("nij,nj->ni", est["node_rotations"], w.nodes) + w.nodes + est["node_translations"]).astype(np.float32)
        ov.set_transforms(est["node_rotations"], T)
        n = ov.integrate(w.frame(f), False)
    assert np.array_equal(vols[1][0], ov.tsdf) and np.array_equal(vols[1][2], ov.weight) and np.array_equal(vols[1][1], ov.colour)
    assert int(counters[1][-1][0, 0]) == n


def test_idempotent_weight_count(oracle):
    """Integrating the same warped frame n times: weight == n exactly on every touched voxel and
    tsdf unchanged to 1 ulp (SURVEY.md 8c known-answer 4)."""
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol = _mk_volume(w)
    wf = WarpField(w.graph(), vol, None)
    vol.warpfield = wf
    vol.fopt.source_frame = -1                       # every frame is a warped frame
    wf.frame_id = -1
    vol.frame_id = -1
    im = w.frame(1)
    est = w.estimate(1)
    first = None
    for f in range(5):
        est["target_frame_id"] = f
        wf.update_transformations(est)
        vol.integrate({"im": im, "id": f})
        t, c, wt = vol.get_volume()
        if first is None:
            first = (t.copy(), c.copy(), wt.copy())
    touched = first[2] > 0
    assert touched.sum() > 1000
    assert np.array_equal(wt[touched], np.full(touched.sum(), 5.0, np.float32))
    assert np.array_equal(wt[~touched], np.zeros((~touched).sum(), np.float32))
    assert np.abs(t - first[0]).max() <= 2.4e-7
    assert np.array_equal(c, first[1])


def test_empty_inputs_and_errors():
    from occlusionfusion_b200 import WarpField, _capi
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    vol = _mk_volume(w)
    wf = WarpField(w.graph(), vol, None)
    a, wt, v = wf.skin(np.zeros((0, 3), np.float32))
    assert a.shape == (0, 4) and wt.shape == (0, 4) and v.shape == (0,)
    vol.update(w.frame(0), 0)
    vis, diff = vol.check_visibility(np.zeros((0, 3), np.float32))
    assert vis.shape == (0,) and diff.shape == (0,)
    # warped frame without a warp field / K > M -> loud errors, no silent fallback
    wf.graph_neighbours = 500
    with pytest.raises(_capi.OfxError):
        wf.skin(np.zeros((4, 3), np.float32))
    with pytest.raises(AssertionError):
        vol.update(w.frame(1), 5)                    # frame id must advance by skip_rate (tsdf.py:569)


def test_full_size_256_properties():
    """BASELINE config c1 (256^3, 1 cm, ~1k nodes): size-independent properties at full size."""
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("c1")
    vol = _mk_volume(w)
    wf = WarpField(w.graph(), vol, None)
    vol.warpfield = wf
    vol.integrate({"im": w.frame(0), "id": 0})
    s0 = vol.last_stats()
    t, c, wt = vol.get_volume()
    assert int(wt.sum()) == s0["updated"] == s0["weight_sum"]
    assert set(np.unique(wt)) <= {0.0, 1.0}
    assert t.min() >= -1.0 and t.max() <= 1.0
    # identity transforms on a warped frame == plain integrate restricted to skin-valid voxels
    est = w.estimate(1)
    est["node_rotations"] = np.tile(np.eye(3, dtype=np.float32), (w.M, 1, 1))
    est["node_translations"] = np.zeros((w.M, 3), np.float32)
    wf.update_transformations(est)
    vol.integrate({"im": w.frame(0), "id": 1})
    s1 = vol.last_stats()
    t1, c1, wt1 = vol.get_volume()
    assert 0 < s1["updated"] <= s1["skin_valid"] < wt.size
    assert int(wt1.sum()) == s0["updated"] + s1["updated"]
    # the warp only ever touches skin-valid voxels, and each at most once per frame
    assert set(np.unique(wt1)) <= {0.0, 1.0, 2.0}
    assert t1.min() >= -1.0 and t1.max() <= 1.0
    assert np.all(c1 < 2 ** 24) and np.array_equal(c1, np.floor(c1))


def _axis_tables(w, oracle):
    """vox2world (tsdf.py:345-348) is separable: one oracle call per axis instead of the full slab."""
    Dx, Dy, Dz = w.dims
    px = oracle.vox2world(w.origin, w.voxel_size, (Dx, 1, 1))[:, 0]
    py = oracle.vox2world(w.origin, w.voxel_size, (1, Dy, 1))[:, 1]
    pz = oracle.vox2world(w.origin, w.voxel_size, (1, 1, Dz))[:, 2]
    return px, py, pz


def _voxel_sample(w, oracle, vol, seed=1234):
    """Seeded sample of voxels of the full-size volume: uniform + around the camera-facing surface (band scenes)
    + whole z-rows.  Returns flat indices (numpy, and as a device tensor) and the voxels' world points."""
    import torch
    Dx, Dy, Dz = w.dims
    rng = np.random.default_rng(seed)
    n_uni = 120_000
    ijk = [np.stack([rng.integers(0, Dx, n_uni), rng.integers(0, Dy, n_uni), rng.integers(0, Dz, n_uni)], 1)]
    if w.scene == "band":
        # voxels around the camera-facing surface of the sphere (where the truncation band and the skin band are)
        n_s = 200_000
        d = rng.normal(size=(n_s, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        d[:, 2] = -np.abs(d[:, 2])
        p = w.centre + (w.radius + rng.uniform(-0.08, 0.08, (n_s, 1))) * d
        q = np.rint((p - w.origin.astype(np.float64)) / w.voxel_size).astype(np.int64)
        ijk.append(q[((q >= 0) & (q < np.array(w.dims))).all(1)])
    rows = np.stack([rng.integers(0, Dx, 48), rng.integers(0, Dy, 48)], 1)          # whole z-rows: every lane of a brick row
    ijk.append(np.concatenate([np.concatenate([np.repeat(r[None], Dz, 0), np.arange(Dz)[:, None]], 1) for r in rows]))
    ijk = np.concatenate(ijk)
    flat = np.unique((ijk[:, 0] * Dy + ijk[:, 1]) * Dz + ijk[:, 2])
    ix, iy, iz = flat // (Dy * Dz), (flat // Dz) % Dy, flat % Dz
    px, py, pz = _axis_tables(w, oracle)
    pts = np.stack([px[ix], py[iy], pz[iz]], 1).astype(np.float32)
    flat_t = torch.from_numpy(flat).to(vol.device)

    return flat, flat_t, pts


@pytest.mark.parametrize("name,n_frames,obs_weight", [("c0", 4, 1.0), ("c1", 4, 1.0), ("c2-dense", 3, 1.0), ("c2", 4, 1.0), ("c4", 3, 2.0)])
def test_full_size_sampled_parity(name, n_frames, obs_weight, oracle):
    """BASELINE configs at FULL size (128^3, 256^3, 512^3 / K=4 as benchmarked; 1024^3, 1280x720, ~4k nodes, K=8): the CUDA path
    integrates the whole volume; a seeded sample of voxels (uniform + concentrated around the surface + whole
    z-rows) is replayed through the oracle's point operators (skin -> warp -> integrate) and must be bit-identical
    after every frame."""
    import torch
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload(name)
    Dx, Dy, Dz = w.dims
    vol = _mk_volume(w)
    wf = WarpField(w.graph(), vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf

    flat, flat_t, pts = _voxel_sample(w, oracle, vol)

    a, wts, valid = oracle.skin(pts, w.nodes, w.K, w.coverage)
    t = np.ones(len(flat), np.float32)
    wt = np.zeros(len(flat), np.float32)
    c = np.zeros(len(flat), np.float32)
    total_updated = 0
    for f in range(n_frames):
        im = w.frame(f)
        color_im, depth_im = oracle.pack_frame(im)
        if f == 0:
            cam_pts, ok = pts, np.ones(len(pts), bool)
        else:
            est = w.estimate(f)
            wf.update_transformations(est)
            cam_pts, ok = oracle.warp(pts, a, wts, valid, est["node_rotations"], wf.translations), valid
        vol.integrate({"im": im, "id": f}, obs_weight)
        oracle.integrate(t, wt, c, cam_pts, ok, depth_im, color_im, w.intr, w.trunc, obs_weight)
        st = vol.last_stats()
        total_updated += st["updated"]
        g_t = vol._tsdf_vol.reshape(-1)[flat_t].cpu().numpy()
        g_w = vol._weight_vol.reshape(-1)[flat_t].cpu().numpy()
        g_c = vol._color_vol.reshape(-1)[flat_t].cpu().numpy()
        assert np.array_equal(g_w, wt), f"frame {f}: weights differ at {(g_w != wt).sum()} of {len(flat)} sampled voxels"
        assert np.array_equal(g_c, c), f"frame {f}: colours differ at {(g_c != c).sum()} sampled voxels"
        assert np.array_equal(g_t, t), f"frame {f}: tsdf differs, max {np.abs(g_t - t).max()}"
    # the sample must have exercised the path: updated voxels, repeated observations, a real surface band
    assert (wt > 0).sum() > 5_000 and wt.max() == n_frames * obs_weight
    if w.scene == "band":
        assert ((t < 1.0) & (t > -1.0)).sum() > 1000 and 0 < valid.sum() < len(valid)
    # whole-volume counter identity at full size: every update adds obs_weight to exactly one voxel
    assert vol._weight_vol.sum(dtype=torch.float64).item() == obs_weight * total_updated


def test_prepare_frame_lookahead_equals_set_transforms():
    """ofx_prepare_frame (node transforms + frustum cull staged ahead, in the other per-frame set) followed by
    ofx_integrate == ofx_set_transforms + ofx_integrate, also when prepared and plain frames alternate; a
    camera that differs from the prepared one is refused."""
    import torch
    from occlusionfusion_b200 import WarpField, _capi
    from occlusionfusion_b200.sharded import CudaShard
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    dev = torch.device("cuda", 0)
    results = []
    for mode in ("plain", "prepared", "mixed"):
        vol = _mk_volume(w)
        wf = WarpField(w.graph(), vol, None)
        vol.warpfield = wf
        be = CudaShard(vol)
        be.set_graph(torch.from_numpy(w.nodes).to(dev), w.coverage, w.K)
        colour, depth = be.pack_frame(w.frame(0))
        be.integrate(depth, colour, True, 1.0)
        stats = []
        for f in range(1, 7):
            est = w.estimate(f)
            T = (-np.einsum("nij,nj->ni", est["node_rotations"], w.nodes) + w.nodes + est["node_translations"]).astype(np.float32)
            R_d, T_d = torch.from_numpy(est["node_rotations"]).to(dev).contiguous(), torch.from_numpy(T).to(dev)
            colour, depth = be.pack_frame(w.frame(f))
            if mode == "prepared" or (mode == "mixed" and f % 3 != 0):
                be.prepare(R_d, T_d, w.H, w.W)
            else:
                be.set_transforms(R_d, T_d)
            stats.append(be.integrate(depth, colour, False, 1.0).cpu().numpy().copy())
        results.append((vol.get_volume(), stats))
        if mode == "mixed":
            be.prepare(R_d, T_d, w.H + 1, w.W)
            with pytest.raises(_capi.OfxError):
                be.integrate(depth, colour, False, 1.0)
    for vols, stats in results[1:]:
        for a, b in zip(results[0][0], vols):
            assert np.array_equal(a, b)
        # updated voxels and weight sums agree frame by frame; bricks / skin-valid voxels VISITED may be larger for a
        # prepared frame (its cull ran before the depth image existed: frustum only, no depth-range test)
        for a, b in zip(results[0][1], stats):
            assert a[0] == b[0] and a[3] == b[3] and b[1] >= a[1] and b[2] >= a[2]
    assert results[0][1][-1][0] > 1000


def test_full_size_sequence_with_node_insertion(oracle):
    """BASELINE config 3 at full size: a 100-frame sequence at 256^3 with per-frame node transforms and node insertion
    every 10th frame (the node table grows 800 -> 1000; old nodes keep their indices, EDGraph.update appends;
    warpfield.py:532-533 re-skins every voxel).  The sampled voxels are replayed through the oracle's point
    operators and compared after every graph change and at the end, bit-identical, weights == number of hits."""
    import torch
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("c1")
    vol = _mk_volume(w)
    nodes = w.nodes[:800].copy()
    wf = WarpField(_graph(nodes, w.coverage), vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf
    flat, flat_t, pts = _voxel_sample(w, oracle, vol, seed=77)
    a, wts, valid = oracle.skin(pts, nodes, w.K, w.coverage)
    t = np.ones(len(flat), np.float32)
    wt = np.zeros(len(flat), np.float32)
    c = np.zeros(len(flat), np.float32)

    def compare(f):
        g_t = vol._tsdf_vol.reshape(-1)[flat_t].cpu().numpy()
        g_w = vol._weight_vol.reshape(-1)[flat_t].cpu().numpy()
        g_c = vol._color_vol.reshape(-1)[flat_t].cpu().numpy()
        assert np.array_equal(g_w, wt), f"frame {f}: weights differ at {(g_w != wt).sum()} of {len(flat)} sampled voxels"
        assert np.array_equal(g_c, c), f"frame {f}: colours differ at {(g_c != c).sum()} sampled voxels"
        assert np.array_equal(g_t, t), f"frame {f}: tsdf differs, max {np.abs(g_t - t).max()}"

    n_valid = [int(valid.sum())]
    for f in range(100):
        im = w.frame(f)
        color_im, depth_im = oracle.pack_frame(im)
        if f == 0:
            cam_pts, ok = pts, np.ones(len(pts), bool)
        else:
            if f % 10 == 0:
                compare(f - 1)
                nodes = w.nodes[: len(nodes) + 20].copy()               # insertion: the table grows, indices are kept
                wf.graph.nodes = nodes
                wf.rotations = np.tile(np.eye(3, dtype=np.float32), (len(nodes), 1, 1))
                wf.translations = np.zeros((len(nodes), 3), np.float32)
                wf.deformed_nodes = nodes.copy()
                a, wts, valid = oracle.skin(pts, nodes, w.K, w.coverage)
                n_valid.append(int(valid.sum()))
            est = w.estimate(f, nodes)
            wf.update_transformations(est)
            cam_pts, ok = oracle.warp(pts, a, wts, valid, est["node_rotations"], wf.translations), valid
        vol.integrate({"im": im, "id": f})
        oracle.integrate(t, wt, c, cam_pts, ok, depth_im, color_im, w.intr, w.trunc, 1.0)
    compare(99)
    assert len(nodes) == 980 and n_valid[-1] > n_valid[0]                 # the inserted nodes reach new voxels
    assert wt.max() >= 30 and ((t < 1.0) & (t > -1.0)).sum() > 1000


@pytest.mark.parametrize("pinned", [False, True])
def test_overlapped_frames_with_a_reused_host_buffer(pinned, oracle):
    """The overlapped end-to-end path (two per-frame sets, upload stream, lazy counters: tsdf.py update / integrate /
    last_stats) under the worst caller: ONE host frame buffer and ONE estimate dict, overwritten for the next frame the
    moment integrate() returns (page-locked buffer: after wait_uploaded(), as update() documents), with nothing else
    waiting for the device.  Every frame differs, so a frame read late, a node-transform staging buffer reused early or
    counters read from the wrong set change the volume or the counters."""
    import torch
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    n_frames = 9
    vol = _mk_volume(w)
    wf = WarpField(w.graph(), vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    ov.set_graph(w.nodes, w.coverage, w.K)
    buf = torch.empty((6, w.H, w.W), dtype=torch.float32, pin_memory=pinned)
    host = buf.numpy()
    est_shared = {k: np.array(v, copy=True) for k, v in w.estimate(1).items() if isinstance(v, np.ndarray)}
    expected, seen = [], []
    for f in range(n_frames):
        im = w.frame(f)
        host[...] = im                                   # overwrite the buffer the previous frame was read from
        if f > 0:
            est = w.estimate(f)
            for k in est_shared:
                est_shared[k][...] = est[k]              # and the arrays the previous transforms were staged from
            wf.update_transformations({**est, **est_shared})
            ov.set_transforms(est["node_rotations"], wf.translations)
        vol.integrate({"im": buf if pinned else host, "id": f})
        if pinned:
            vol.wait_uploaded()
        host[...] = np.float32(7.5)                      # garbage, at once
        for k in est_shared:
            est_shared[k][...] = 3.0
        expected.append(ov.integrate(im, f == 0))
        prev = vol.last_stats(previous=True)             # never waits for the frame just submitted
        if f > 0:
            assert prev is not None
            seen.append(prev["updated"])
    seen.append(vol.last_stats()["updated"])
    assert seen == expected and min(expected[1:]) > 1000
    _assert_bit_equal(vol, ov)


@pytest.mark.parametrize("cap", ["1", "6", "32"])
def test_row_table_capacity_does_not_change_results(cap, oracle, monkeypatch):
    """The persistent kernel sizes its per-warp row tables per graph (choose_tab_cap, csrc/ofx_skin_cache.cu); bricks with
    more nodes than the capacity blend straight from the staged node transforms.  Forced capacities (the library's
    investigation hook OFX_TAB_CAP): 1 = every brick takes the staged-transform path, 6 = a mix, 32 = every brick the
    tables -- all bit-identical to the oracle."""
    from occlusionfusion_b200.synthetic import Workload
    monkeypatch.setenv("OFX_TAB_CAP", cap)
    w = Workload("tiny")
    vol, ov, wf, stats = _run_pair(w, oracle, n_frames=4)
    _assert_bit_equal(vol, ov)
    for st, n in stats:
        assert st["updated"] == n


@pytest.mark.parametrize("pinned", [False, True])
def test_pipelined_driver_on_host_frames(pinned, oracle):
    """The end-to-end form of the pipelined driver (what bench.py's e2e leg runs at N > 1): pipeline_step is handed HOST
    frames + node transforms as (im, R, T) tuples, stages them on the communication stream (sharded.py _stage_host: DMA of
    RGB + Z, pack kernel, R|T through pinned memory) and reads the counters one exchange late from pinned memory.  One
    shard, bit-identical to the oracle; the host buffers are overwritten as soon as the documented rule allows."""
    import torch
    from occlusionfusion_b200 import WarpField
    from occlusionfusion_b200.sharded import CudaShard, ShardedFusion
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    dev = torch.device("cuda", 0)
    vol = _mk_volume(w)
    wf = WarpField(w.graph(), vol, None)
    vol.warpfield = wf
    sf = ShardedFusion(CudaShard(vol), w.dims, 0, 1, w.dims[0], dev)
    sf.set_graph(w.nodes, w.coverage, w.K)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    ov.set_graph(w.nodes, w.coverage, w.K)
    sf.fill_packet(w.frame(0), np.tile(np.eye(3, dtype=np.float32), (w.M, 1, 1)), np.zeros((w.M, 3), np.float32))
    sf.step(w.H, w.W, identity_warp=True)
    ov.integrate(w.frame(0), True)
    n = 7
    bufs = [torch.empty((6, w.H, w.W), dtype=torch.float32, pin_memory=pinned) for _ in range(2)]   # two in flight at most

    def host_packet(i):
        f = i + 1
        est = w.estimate(f)
        T = (-np.einsum("nij,nj->ni", est["node_rotations"], w.nodes) + w.nodes + est["node_translations"]).astype(np.float32)
        b = bufs[i % 2]
        b.numpy()[...] = w.frame(f)
        return (b.numpy(), est["node_rotations"], T)

    expected = []
    for i in range(n):
        f = i + 1
        est = w.estimate(f)
        T = (-np.einsum("nij,nj->ni", est["node_rotations"], w.nodes) + w.nodes + est["node_translations"]).astype(np.float32)
        ov.set_transforms(est["node_rotations"], T)
        expected.append(ov.integrate(w.frame(f), False))
    sf.pipeline_begin(w.H, w.W, host_packet(0))
    for i in range(n):
        if pinned and i >= 1:
            sf.wait_staged(i - 1)            # the buffer packet i+1 is about to be written into was packet i-1's
        sf.pipeline_step(host_packet(i + 1) if i + 1 < n else None, last=(i + 1 == n))
    sf.pipeline_end()
    got = [int(sf.counters_host(t)[:, 0].sum()) for t in range(n - 2, n)]
    assert got == expected[-2:] and min(expected) > 1000
    _assert_bit_equal(vol, ov)
