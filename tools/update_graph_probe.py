"""Investigation only: where WarpField.update_graph() spends its time (c1, 256^3): wraps the stages in synchronised timers.
    python tools/update_graph_probe.py"""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from occlusionfusion_b200 import EDGraph, TSDFVolume, WarpField, embedded_deformation_graph as edg, graph_proc
from occlusionfusion_b200.synthetic import Workload

T = {}
def wrap(obj, name, label=None):
    fn = getattr(obj, name)
    label = label or name
    def inner(*a, **k):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = fn(*a, **k)
        torch.cuda.synchronize(); T.setdefault(label, []).append((time.perf_counter() - t0) * 1e3)
        return r
    setattr(obj, name, inner)

w = Workload(os.environ.get("WORKLOAD", "c1"))
vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
frames = [torch.from_numpy(w.frame(f)).pin_memory().numpy() for f in range(8)]
vol.integrate({"im": frames[0], "id": 0})
graph = EDGraph(tsdf=vol, nodes=np.zeros((0, 3), np.float32), graph_generation_parameters={"node_coverage": w.coverage})
graph.create_graph_from_mesh(*vol.get_canonical_model()[:2])
wf = WarpField(graph, vol, None); wf.graph_neighbours = w.K; vol.warpfield = wf
for name in ("get_canonical_model", "get_truncated_region"):
    wrap(vol, name)
_mc = TSDFVolume.marching_cubes
def _mc_timed(*a, **k):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    r = _mc(*a, **k)
    torch.cuda.synchronize(); T.setdefault("marching_cubes", []).append((time.perf_counter() - t0) * 1e3)
    return r
vol.marching_cubes = _mc_timed
for name in ("erode_mesh", "update", "rebuild_edges", "remove_nodes_with_not_enough_neighbours", "compute_clusters", "create_graph_pyramid", "get_reduced_graph"):
    wrap(graph, name)
wrap(wf, "find_unreachable_nodes"); wrap(wf, "_sync_device"); wrap(wf, "update_graph")
wrap(edg, "compute_edges_geodesic_d", "geodesic_edges(device)"); wrap(edg, "downsample_nodes_d", "downsample_nodes(device)"); wrap(edg, "DeviceMesh", "DeviceMesh upload")
for f in range(1, 41):
    if f % 10 == 0:
        vol.clear(); wf.update_graph()
    est = w.estimate(f, wf.graph.nodes)
    wf.update_transformations(est)
    vol.integrate({"im": frames[f % 8], "id": f})
verts, faces = vol.get_canonical_model()[:2]
print(f"mesh {len(verts)} vertices {len(faces)} faces, {wf.graph.nodes.shape[0]} nodes")
for k, v in T.items():
    print(f"  {k:45s} n={len(v):3d} median {np.median(v):8.3f} ms  total/update_graph {sum(v) / max(1, len(T['update_graph'])):8.3f} ms")
if os.environ.get("PROBE_LISTS"):      # every call, in order: the first calls carry one-off costs
    for k, v in T.items():
        print(f"  {k:45s} " + " ".join(f"{x:.2f}" for x in v))
