"""Investigation only: where an end-to-end frame (public classes, host buffers) spends its time.  tools/e2e_probe.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from occlusionfusion_b200 import TSDFVolume, WarpField
from occlusionfusion_b200.synthetic import Workload

w = Workload(os.environ.get("WORKLOAD", "c2-dense"))
dev = torch.device("cuda:0")
vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None, device=dev)
wf = WarpField(w.graph(), vol, None)
vol.warpfield = wf
frames = [torch.from_numpy(w.frame(i)).pin_memory().numpy() for i in range(5)]
ests = [w.estimate(i) for i in range(1, 5)]
vol.integrate({"im": frames[0], "id": 0})
acc = {}
def lap(name, t0, sync=False):
    if sync:
        torch.cuda.synchronize()
    t = time.perf_counter()
    acc.setdefault(name, []).append((t - t0) * 1e6)
    return t
for mode in ("async", "sync-each"):
    acc.clear()
    sy = mode == "sync-each"
    for s in range(1, 40):
        est = dict(ests[s % 4]); est["target_frame_id"] = vol.frame_id + 1
        wf.frame_id = vol.frame_id
        torch.cuda.synchronize()
        t0 = ta = time.perf_counter()
        wf.update_transformations(est); t0 = lap("update_transformations (host numpy)", t0, sy)
        vol.update(frames[1 + s % 4], vol.frame_id + 1); t0 = lap("update: H2D RGB+Z, pack kernel", t0, sy)
        wf._sync_device(); t0 = lap("H2D R|T + node-transform kernel", t0, sy)
        H, W = vol._depth_im_dev.shape
        from occlusionfusion_b200 import _capi
        _capi.check(vol._lib.ofx_integrate(vol._h, _capi.ptr(vol._depth_im_dev), _capi.ptr(vol._color_im_dev), H, W, vol._intr4, 1.0, 0,
                                           _capi.ptr(vol._stats), _capi.stream_ptr(dev))); t0 = lap("ofx_integrate (cull + fused)", t0, sy)
        vol.last_stats(); t0 = lap("last_stats (D2H + sync)", t0)
        acc.setdefault("TOTAL", []).append((t0 - ta) * 1e6)
    print(mode)
    for k, v in acc.items():
        print(f"   {k:42s} {np.median(v[5:]):8.1f} us")
