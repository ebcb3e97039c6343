"""Investigation only: where the HOST time of the single-GPU public frame loop goes (c1: the device needs 0.03 ms per frame,
so the loop is host-paced).  cProfile over N frames of update_transformations + integrate + last_stats(previous=True).
    python tools/host_profile.py [frames]"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np                                                    # noqa: E402
import torch                                                          # noqa: E402
from occlusionfusion_b200 import TSDFVolume, WarpField                # noqa: E402
from occlusionfusion_b200.synthetic import Workload                   # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
w = Workload(os.environ.get("WORKLOAD", "c1"))
vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
wf = WarpField(w.graph(), vol, None)
wf.graph_neighbours = w.K
vol.warpfield = wf
frames = [torch.from_numpy(w.frame(f)).pin_memory().numpy() for f in range(5)]
ests = [w.estimate(f) for f in range(1, 5)]
vol.integrate({"im": frames[0], "id": 0})


def loop(k, start):
    upd = 0
    for i in range(k):
        est = dict(ests[i % 4])
        est["target_frame_id"] = start + i
        wf.update_transformations(est)
        vol.integrate({"im": frames[1 + i % 4], "id": start + i})
        if i:
            upd += vol.last_stats(previous=True)["updated"]
    return upd


loop(50, 1)
torch.cuda.synchronize()
t0 = time.perf_counter()
loop(n, 51)
torch.cuda.synchronize()
print(f"{w.name}: {(time.perf_counter() - t0) * 1e6 / n:.1f} us per frame (host wall clock, {n} frames)")
pr = cProfile.Profile()
pr.enable()
loop(n, 51 + n)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
