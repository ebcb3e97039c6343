"""BASELINE config 3 in its own shape: a frame sequence at 256^3 with a per-frame graph-node update and node insertion every
10th frame, all through the drop-in classes the way fusion_with_occlusion/fusion.py:202-221 drives them --
    warpfield.update_transformations(estimate); tsdf.integrate(frame); every 10th frame: tsdf.clear(); warpfield.update_graph()
-- with the canonical mesh taken from the volume itself (truncated region -> marching cubes), erosion, unreachable vertices,
greedy insertion, edge rebuild and re-skinning.  Prints one JSON object with the time per phase (host wall clock around
synchronised calls; the estimator that would produce the transforms is outside the path and outside the timing).
    python tools/sequence_c3.py [--frames 100] > gpurun_out/sequence_c3.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from occlusionfusion_b200 import EDGraph, TSDFVolume, WarpField  # noqa: E402
from occlusionfusion_b200.synthetic import Workload  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=100)
ap.add_argument("--workload", default="c1")
args = ap.parse_args()

w = Workload(args.workload)
vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
frames = [torch.from_numpy(w.frame(f)).pin_memory().numpy() for f in range(8)]


def timed(fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3, r


t_src, _ = timed(lambda: vol.integrate({"im": frames[0], "id": 0}))
# the graph comes from the source frame's own surface, as in fusion.py:74 (EDGraph(tsdf, vis, source_data))
graph = EDGraph(tsdf=vol, nodes=np.zeros((0, 3), np.float32), graph_generation_parameters={"node_coverage": w.coverage})
t_build, _ = timed(lambda: graph.create_graph_from_mesh(*vol.get_canonical_model()[:2]))
nodes = graph.nodes
wf = WarpField(graph, vol, None)
wf.graph_neighbours = w.K
vol.warpfield = wf
t_int, t_upd, t_graph, growth, updated = [], [], [], [len(nodes)], []
for f in range(1, args.frames):
    if f % 10 == 0:
        vol.clear()
        ms, changed = timed(lambda: wf.update_graph())
        t_graph.append(ms)
        growth.append(int(wf.graph.nodes.shape[0]))
    est = w.estimate(f, wf.graph.nodes)                     # stands in for the estimator (not timed)
    ms, _ = timed(lambda: wf.update_transformations(est))
    t_upd.append(ms)
    ms, _ = timed(lambda: vol.integrate({"im": frames[f % len(frames)], "id": f}))
    t_int.append(ms)
    updated.append(vol.last_stats()["updated"])
verts, faces, _, _ = vol.get_canonical_model()
print(json.dumps({
    "workload": f"{w.name}: {w.dims} voxels @ {w.voxel_size * 1e3:.1f} mm, {w.W}x{w.H}, K={w.K}, {args.frames} frames, node insertion every 10th frame",
    "gpu": torch.cuda.get_device_name(0), "frames": args.frames,
    "source_frame_ms": t_src, "graph_from_source_mesh_ms": t_build,
    "integrate_ms_median": float(np.median(t_int)), "integrate_ms_mean": float(np.mean(t_int)),
    "update_transformations_ms_median": float(np.median(t_upd)),
    "update_graph_ms": [round(x, 2) for x in t_graph], "nodes": growth,
    "updated_voxels_per_frame_median": float(np.median(updated)),
    "ms_per_frame_all_in": float((sum(t_int) + sum(t_upd) + sum(t_graph)) / max(1, args.frames - 1)),
    "final_mesh": {"vertices": int(len(verts)), "faces": int(len(faces))},
    "timing": "host wall clock around synchronised calls: host frames in pinned memory, H2D inside integrate",
}))
