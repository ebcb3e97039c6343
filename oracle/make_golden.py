"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by RUNNING THE REFERENCE'S OWN
PYTHON (CPU branch) in this container.  Usage:  python oracle/make_golden.py

For each case it drives the reference objects exactly as fusion.py:57-83,176-179 does:
    TSDFVolume(bbox, max_depth, cam_intr, fopt, None)           tsdf.py:41
    tsdf.integrate(frame 0)      (source frame, identity warp)  tsdf.py:378, 395-398
    WarpField(graph, tsdf, None)                                warpfield.py:22
    warpfield.update_transformations(estimate_f)                warpfield.py:389
    tsdf.integrate(frame f)      (warped frame, CPU branch)     tsdf.py:401, 442-494
and stores inputs and every intermediate the parity tests compare against.

`WarpField.deform` is hard-wired to `self.optimizer.deform_ED` (warpfield.py:75,274), which
needs lietorch (missing submodule).  The stand-in optimizer below keeps deform_ED's contract
(registration_fusion.py:157-184: invalid points pass through unchanged) and evaluates the
blend with the reference's own numba `WarpField.deform_lbs` (warpfield.py:208-231) on the
valid subset -- calling it on the subset avoids the whole-array assignment bug at
warpfield.py:229.  A float64 evaluation of the ED form (geometry.py:23-24) is stored next to
it as `warped_ed64_*` for reference.
"""
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(HERE), "tests", "golden")


def rodrigues(rv):
    """rotation vectors (M,3) -> matrices (M,3,3), float64."""
    th = np.linalg.norm(rv, axis=1)
    k = rv / np.maximum(th, 1e-12)[:, None]
    K = np.zeros((len(rv), 3, 3))
    K[:, 0, 1], K[:, 0, 2] = -k[:, 2], k[:, 1]
    K[:, 1, 0], K[:, 1, 2] = k[:, 2], -k[:, 0]
    K[:, 2, 0], K[:, 2, 1] = -k[:, 1], k[:, 0]
    s, c = np.sin(th)[:, None, None], np.cos(th)[:, None, None]
    return np.eye(3)[None] + s * K + (1 - c) * (K @ K)


def sphere_frame(H, W, intr, centre, radius, seed):
    """(6,H,W) float32 frame: RGB in [0,1] + XYZ metres (layout model/dataset.py:202-206)."""
    fx, fy, cx, cy = intr
    v, u = np.mgrid[0:H, 0:W].astype(np.float64)
    d = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], 0)
    a = (d * d).sum(0)
    b = -2 * (d * np.asarray(centre)[:, None, None]).sum(0)
    c = float(np.dot(centre, centre)) - radius * radius
    disc = b * b - 4 * a * c
    t = np.where(disc > 0, (-b - np.sqrt(np.maximum(disc, 0))) / (2 * a), 0.0)
    im = np.zeros((6, H, W), np.float32)
    rng = np.random.default_rng(seed)
    im[:3] = rng.integers(0, 256, (3, H, W)).astype(np.float32) / 255.0
    im[3:] = (d * t).astype(np.float32)
    return im


class LbsOptimizer:
    """Stand-in for PytorchRegistration on the one method WarpField.deform calls."""

    def __init__(self, warpfield, ref_warpfield_cls):
        self.wf, self.cls = warpfield, ref_warpfield_cls

    def deform_ED(self, points, anchors, weights, valid_pts):
        out = points.copy()                                   # registration_fusion.py:161
        idx = np.nonzero(valid_pts)[0]
        if len(idx):
            sub = self.cls.deform_lbs(self.wf.rotations.astype(np.float32),
                                      self.wf.translations.astype(np.float32),
                                      np.ascontiguousarray(points[idx]),
                                      np.ascontiguousarray(anchors[idx]),
                                      np.ascontiguousarray(weights[idx]),
                                      np.ones(len(idx), bool))
            out[idx] = sub                                    # registration_fusion.py:180
        return out


def run_case(name, *, H, W, voxel_dim, bbox, max_depth, n_nodes, coverage, K, n_frames, seed,
             radius=0.35):
    ref_tsdf, ref_wf = ref_shims.import_reference()
    rng = np.random.default_rng(seed)
    intr = np.array([525.0 * W / 640, 525.0 * W / 640, (W - 1) / 2, (H - 1) / 2])
    tmp = tempfile.mkdtemp(prefix="ofx_golden_")
    fopt = types.SimpleNamespace(voxel_dim=voxel_dim, source_frame=0, skip_rate=1,
                                 datadir=tmp, gpu=False)
    centre = np.array([0.0, 0.0, 0.62 * max_depth])
    frames = [sphere_frame(H, W, intr, centre + 0.01 * f, radius, seed + f) for f in range(n_frames)]

    vol = ref_tsdf.TSDFVolume(bbox, max_depth, intr, fopt, None)
    out = dict(H=H, W=W, intr=intr, bbox=np.array(bbox, np.float64), max_depth=max_depth,
               voxel_dim=np.array(vol._vol_dim), K=K, coverage=coverage,
               vol_origin=vol._vol_origin, voxel_size=np.float64(vol._voxel_size),
               trunc=np.float64(vol._trunc_margin), frames=np.stack(frames),
               world_pts=vol.world_pts)
    vol.integrate({"im": frames[0], "id": 0})
    out["color_im_0"], out["depth_im_0"] = vol.color_im, vol.depth_im
    t, c, w = vol.get_volume() if hasattr(vol, "_tsdf_vol_gpu") else (
        vol._tsdf_vol_cpu, vol._color_vol_cpu, vol._weight_vol_cpu)
    out["tsdf_0"], out["color_0"], out["weight_0"] = t.copy(), c.copy(), w.copy()

    # graph nodes: points on the sphere facing the camera, in general position
    dirs = rng.normal(size=(n_nodes, 3))
    dirs[:, 2] = -np.abs(dirs[:, 2])
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    nodes = (centre + radius * dirs + rng.uniform(-1e-3, 1e-3, (n_nodes, 3))).astype(np.float32)
    graph = types.SimpleNamespace(nodes=nodes,
                                  graph_generation_parameters={"node_coverage": coverage})
    wf = ref_wf.WarpField(graph, vol, None)
    wf.graph_neighbours = min(n_nodes, K)
    wf.optimizer = LbsOptimizer(wf, ref_wf.WarpField)
    vol.warpfield = wf
    out["nodes"] = nodes

    for f in range(1, n_frames):
        rv = 0.15 * np.sin(2 * np.pi * nodes.astype(np.float64) / 1.3 + 0.05 * f)
        R = rodrigues(rv).astype(np.float32)
        tr = (0.02 * np.cos(2 * np.pi * nodes.astype(np.float64) / 0.9 + 0.05 * f)).astype(np.float32)
        est = dict(node_rotations=R, node_translations=tr,
                   deformed_nodes_to_target=nodes + tr, source_frame_id=0, target_frame_id=f)
        wf.update_transformations(est)
        vol.integrate({"im": frames[f], "id": f})
        warped, valid = wf.deform_tsdf()   # same state integrate() just used (cached skin)
        out[f"R_{f}"], out[f"t_{f}"] = R, tr
        out[f"T_{f}"] = np.asarray(wf.translations)
        out[f"warped_{f}"], out[f"valid_{f}"] = warped, valid
        g = nodes.astype(np.float64)
        a = np.where(wf.world_anchors >= 0, wf.world_anchors, 0)
        x = vol.world_pts.astype(np.float64)[:, None, :]
        ed = (np.einsum("nkij,nkj->nki", R.astype(np.float64)[a], x - g[a]) + g[a]
              + tr.astype(np.float64)[a]) * wf.world_weights.astype(np.float64)[..., None]
        out[f"warped_ed64_{f}"] = np.where(valid[:, None], ed.sum(1), vol.world_pts)
        out[f"tsdf_{f}"] = vol._tsdf_vol_cpu.copy()
        out[f"color_{f}"] = vol._color_vol_cpu.copy()
        out[f"weight_{f}"] = vol._weight_vol_cpu.copy()
        out[f"color_im_{f}"], out[f"depth_im_{f}"] = vol.color_im, vol.depth_im
    out["anchors"], out["weights"] = wf.world_anchors, wf.world_weights
    out["skin_valid"] = wf.world_valid_pts

    # free-standing skin() on random points against an explicit node set (warpfield.py:99-104)
    q = (centre + rng.uniform(-0.6, 0.6, (2000, 3))).astype(np.float32)
    a, w_, v = wf.skin(q, nodes=nodes[: max(K, n_nodes // 2)])
    out["skin_q"], out["skin_q_anchors"], out["skin_q_weights"], out["skin_q_valid"] = q, a, w_, v
    # check_visibility on the same points against the last frame (tsdf.py:599-612)
    vis_ok, vis_diff = vol.check_visibility(q)
    out["vis_valid"], out["vis_diff"] = vis_ok, vis_diff

    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, f"{name}.npz")
    np.savez_compressed(path, **out)
    n_upd = [(out[f"weight_{f}"] != out[f"weight_{f-1}"]).sum() for f in range(1, n_frames)]
    print(f"{name}: dims={out['voxel_dim']} vs={out['voxel_size']:.6f} M={n_nodes} K={K} "
          f"skin-valid={int(out['skin_valid'].sum())} updated/frame={n_upd} "
          f"-> {path} ({os.path.getsize(path)/1e3:.0f} kB)")


if __name__ == "__main__":
    run_case("ref_k4", H=48, W=64, voxel_dim=[28, 32, 36], bbox=(4, 2, 60, 46), max_depth=1.6,
             n_nodes=60, coverage=0.15, K=4, n_frames=3, seed=20221)
    run_case("ref_k8", H=60, W=80, voxel_dim=24, bbox=(0, 0, 80, 60), max_depth=1.5,
             n_nodes=64, coverage=0.14, K=8, n_frames=2, seed=20222)
    run_case("ref_m3", H=48, W=64, voxel_dim=16, bbox=(0, 0, 64, 48), max_depth=1.6,
             n_nodes=3, coverage=0.2, K=4, n_frames=2, seed=20223)
