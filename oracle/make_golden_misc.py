"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/ref_misc.npz by RUNNING THE REFERENCE'S OWN PYTHON in this
container (shims: oracle/ref_shims.py).  Usage:  python oracle/make_golden_misc.py

Pins the corners of the drop-in classes that the sequence goldens (make_golden.py) do not reach:
    TSDFVolume.save_volume / load_volume            tsdf.py:682-702   the bytes of a pickle written by the reference
    TSDFVolume.get_visible_nodes                    tsdf.py:614-638   result + the visible_nodes/<id>.npy file it writes
    WarpField.update_transformations' side file     warpfield.py:418  deformed_nodes/<id>.npy
    WarpField.deform_normals                        warpfield.py:312-345 (CPU branch: numba deform_lbs, translations = 0)
    WarpField.get_transformation_wrt_graph_node / _wrt_origin   warpfield.py:422-449
"""
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402
from make_golden import GOLDEN_DIR, LbsOptimizer, rodrigues, sphere_frame  # noqa: E402


def main():
    ref_tsdf, ref_wf = ref_shims.import_reference()
    H, W, K, n_nodes, coverage, max_depth, radius, seed = 48, 64, 4, 40, 0.2, 1.6, 0.35, 20231
    rng = np.random.default_rng(seed)
    intr = np.array([525.0 * W / 640, 525.0 * W / 640, (W - 1) / 2, (H - 1) / 2])
    tmp = tempfile.mkdtemp(prefix="ofx_golden_misc_")
    fopt = types.SimpleNamespace(voxel_dim=[16, 20, 24], source_frame=0, skip_rate=1, datadir=tmp, gpu=False)
    centre = np.array([0.0, 0.0, 0.62 * max_depth])
    frames = [sphere_frame(H, W, intr, centre + 0.01 * f, radius, seed + f) for f in range(2)]
    bbox = (0, 0, W, H)
    vol = ref_tsdf.TSDFVolume(bbox, max_depth, intr, fopt, None)
    vol.integrate({"im": frames[0], "id": 0})
    dirs = rng.normal(size=(n_nodes, 3))
    dirs[:, 2] = -np.abs(dirs[:, 2])
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    nodes = (centre + radius * dirs + rng.uniform(-1e-3, 1e-3, (n_nodes, 3))).astype(np.float32)
    graph = types.SimpleNamespace(nodes=nodes, graph_generation_parameters={"node_coverage": coverage})
    wf = ref_wf.WarpField(graph, vol, None)
    wf.graph_neighbours = K
    wf.gpu = False
    wf.optimizer = LbsOptimizer(wf, ref_wf.WarpField)
    vol.warpfield = wf
    rv = 0.15 * np.sin(2 * np.pi * nodes.astype(np.float64) / 1.3 + 0.05)
    R = rodrigues(rv).astype(np.float32)
    tr = (0.02 * np.cos(2 * np.pi * nodes.astype(np.float64) / 0.9 + 0.05)).astype(np.float32)
    est = dict(node_rotations=R, node_translations=tr, deformed_nodes_to_target=nodes + tr, source_frame_id=0,
               target_frame_id=1)
    out = dict(H=H, W=W, intr=intr, bbox=np.array(bbox, np.float64), max_depth=max_depth, voxel_dim=np.array(vol._vol_dim),
               K=K, coverage=coverage, frames=np.stack(frames), nodes=nodes, R=R, t=tr)
    wf.update_transformations(est)
    out["T"] = np.asarray(wf.translations)
    out["deformed_nodes_npy"] = np.frombuffer(open(os.path.join(tmp, "results", "deformed_nodes", "1.npy"), "rb").read(), np.uint8)
    Rn, tn = wf.get_transformation_wrt_graph_node()
    out["R_wrt_node"], out["t_wrt_node"] = np.asarray(Rn), np.asarray(tn)
    Ro, To = wf.get_transformation_wrt_origin(Rn, tn)
    out["R_wrt_origin"], out["T_wrt_origin"] = np.asarray(Ro), np.asarray(To)
    vol.integrate({"im": frames[1], "id": 1})

    # visibility of the deformed nodes in frame 1 (tsdf.py:614-638) and the side file
    vis = vol.get_visible_nodes()
    out["visible_nodes"] = np.asarray(vis)
    out["visible_nodes_npy"] = np.frombuffer(open(os.path.join(tmp, "results", "visible_nodes", "1.npy"), "rb").read(), np.uint8)

    # deform_normals on points all of whose K anchors are in range (the reference's numba deform_lbs cannot take a mixed
    # mask, warpfield.py:229): points next to the nodes, random unit normals
    pts = (nodes[rng.integers(0, n_nodes, 400)] + rng.uniform(-0.03, 0.03, (400, 3))).astype(np.float32)
    a, w_, v = wf.skin(pts)
    keep = v
    pts, a, w_ = np.ascontiguousarray(pts[keep]), np.ascontiguousarray(a[keep]), np.ascontiguousarray(w_[keep])
    nrm = rng.normal(size=pts.shape)
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    dn = wf.deform_normals(nrm, a.astype(np.int32), w_, [len(pts), 1, 1], np.ones(len(pts), bool))
    out["dn_points"], out["dn_normals"], out["dn_anchors"], out["dn_weights"] = pts, nrm, a.astype(np.int32), w_
    out["dn_out"] = np.asarray(dn)

    # the volume pickle (tsdf.py:682-687)
    pk = os.path.join(tmp, "volume.pkl")
    vol.save_volume(pk)
    out["volume_pickle"] = np.frombuffer(open(pk, "rb").read(), np.uint8)
    out["tsdf"], out["colour"], out["weight"] = vol._tsdf_vol_cpu.copy(), vol._color_vol_cpu.copy(), vol._weight_vol_cpu.copy()
    vol2 = ref_tsdf.TSDFVolume(bbox, max_depth, intr, fopt, None)
    vol2.load_volume(pk)
    assert np.array_equal(vol2._tsdf_vol_cpu, out["tsdf"]) and np.array_equal(vol2._weight_vol_cpu, out["weight"])

    path = os.path.join(GOLDEN_DIR, "ref_misc.npz")
    np.savez_compressed(path, **out)
    print(f"ref_misc: dims={out['voxel_dim']} visible {int(out['visible_nodes'].sum())}/{n_nodes} normals {len(pts)} "
          f"pickle {len(out['volume_pickle'])} B -> {path} ({os.path.getsize(path) / 1e3:.0f} kB)")


if __name__ == "__main__":
    main()
