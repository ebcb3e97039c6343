"""Times the graph-maintenance / surface-mask operators of libofx.so (ofx_graph.cu) with CUDA events and prints one JSON
object: nearest node, erode_mesh, sample_nodes, greedy insertion on a 300x300 / 700x700 height-field mesh, and the
truncated region at 256^3 / 512^3 (5 B/voxel algorithmic: 4 read + 1 written) against the measured HBM peak.
    python tools/graph_probe.py > gpurun_out/graph_probe.json
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from occlusionfusion_b200 import _capi  # noqa: E402

lib = _capi.load()
dev = torch.device("cuda", 0)
S = _capi.stream_ptr


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def mesh(n, rng):
    u, v = np.meshgrid(np.linspace(-0.8, 0.8, n), np.linspace(-0.6, 0.6, n), indexing="ij")
    z = 1.0 + 0.2 * np.sin(4 * u) * np.cos(3 * v)
    verts = (np.stack([u, v, z], -1).reshape(-1, 3) + rng.uniform(-1e-3, 1e-3, (n * n, 3))).astype(np.float32)
    ij = np.argwhere(rng.random((n - 1, n - 1)) > 0.05)
    a = ij[:, 0] * n + ij[:, 1]
    faces = np.concatenate([np.stack([a, a + 1, a + n], 1), np.stack([a + 1, a + n + 1, a + n], 1)]).astype(np.int32)
    return verts, faces


out = {"gpu": torch.cuda.get_device_name(0)}
rng = np.random.default_rng(11)
for n in (300, 700):
    verts, faces = mesh(n, rng)
    V, F = len(verts), len(faces)
    v_d, f_d = torch.from_numpy(verts).to(dev), torch.from_numpy(faces).to(dev)
    ne = torch.empty(V, dtype=torch.uint8, device=dev)
    t_erode = timed(lambda: _capi.check(lib.ofx_erode_mesh(_capi.ptr(f_d), F, V, 3, 3, _capi.ptr(ne), S())))
    idx = torch.empty(V, dtype=torch.int32, device=dev)
    pos = torch.empty((V, 3), dtype=torch.float32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    import ctypes
    t_sample = timed(lambda: _capi.check(lib.ofx_sample_nodes(_capi.ptr(v_d), V, _capi.ptr(ne), ctypes.c_float(0.05), 1,
                                                              _capi.ptr(idx), _capi.ptr(pos), _capi.ptr(cnt), S())), reps=3)
    M = int(cnt.item())
    nodes = pos[: M // 2].contiguous()                       # a graph that covers about half of the mesh
    dist = torch.empty(V, dtype=torch.float32, device=dev)
    t_near = timed(lambda: _capi.check(lib.ofx_nearest_node(_capi.ptr(v_d), V, _capi.ptr(nodes), nodes.shape[0], _capi.ptr(dist),
                                                            None, S())))
    d = dist.cpu().numpy()
    cand = np.where(d > 0.1)[0]
    cand = cand[np.argsort(d[cand])[::-1]].astype(np.int32)
    c_d = torch.from_numpy(cand).to(dev)
    acc = torch.empty(max(len(cand), 1), dtype=torch.int32, device=dev)
    t_greedy = timed(lambda: _capi.check(lib.ofx_greedy_insert(_capi.ptr(v_d), V, _capi.ptr(c_d), len(cand), _capi.ptr(nodes),
                                                               nodes.shape[0], 0.05, _capi.ptr(acc), _capi.ptr(cnt), S())), reps=3)
    out[f"mesh_{n}"] = dict(V=V, F=F, erode3_ms=t_erode, sample_nodes_ms=t_sample, nodes=M, nearest_ms=t_near,
                            nearest_gpairs_per_s=V * nodes.shape[0] / t_near / 1e6, candidates=len(cand),
                            greedy_insert_ms=t_greedy, inserted=int(cnt.item()))

peaks = {}
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    peaks = json.load(open(p))
for D in (256, 512):
    g = torch.Generator(device=dev).manual_seed(3)
    z = torch.linspace(-6, 6, D, device=dev).view(1, 1, D).expand(D, D, D)
    t = (z + 0.3 * torch.randn((D, D, 1), device=dev, generator=g)).clamp(-1, 1).contiguous()   # a wavy surface: ~8 % band
    reg = torch.empty((D, D, D), dtype=torch.uint8, device=dev)
    ms = timed(lambda: _capi.check(lib.ofx_truncated_region(_capi.ptr(t), D, D, D, 1.2, _capi.ptr(reg), S())))
    gbs = 5.0 * D ** 3 / ms / 1e6
    out[f"truncated_region_{D}"] = dict(ms=ms, band_frac=float((t.abs() <= 0.9).float().mean()), region_frac=float(reg.float().mean()),
                                        achieved_gbs=gbs, peak_gbs=peaks.get("hbm_gbs"),
                                        frac=(gbs / peaks["hbm_gbs"]) if peaks.get("hbm_gbs") else None)
print(json.dumps(out))
