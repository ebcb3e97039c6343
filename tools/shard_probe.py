"""Investigation only: one GPU plays shard 0 of N (c2-dense) and reports the per-frame device times of the
node-transform kernel, the cull and the fused kernel, plus (OFX_TIMING build) the per-CTA timeline.
  tools/shard_probe.py N [name of an OFX_TIMING build]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
timing = len(sys.argv) > 2          # name of an OFX_TIMING build: libofx_<name>.so
if timing:
    os.environ["OFX_LIB"] = os.path.join(ROOT, "occlusionfusion_b200", "lib", f"libofx_{sys.argv[2]}.so")
import numpy as np
import torch
from occlusionfusion_b200 import TSDFVolume, WarpField, _capi
from occlusionfusion_b200.sharded import CudaShard, default_x_block
from occlusionfusion_b200.synthetic import Workload

w = Workload(os.environ.get("WORKLOAD", "c2-dense"))
dev = torch.device("cuda:0")
xb = default_x_block(w.dims[0], N)
vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None, device=dev, shard=0, n_shards=N, x_block=xb)
wf = WarpField(w.graph(), vol, None)
vol.warpfield = wf
be = CudaShard(vol)
be.set_graph(torch.from_numpy(w.nodes).to(dev), w.coverage, w.K)
frames = []
for i in range(1, 5):
    e = w.estimate(i)
    T = (-np.einsum("nij,nj->ni", e["node_rotations"], w.nodes) + w.nodes + e["node_translations"]).astype(np.float32)
    colour, depth = be.pack_frame(w.frame(i))
    frames.append((torch.from_numpy(e["node_rotations"]).to(dev).contiguous(), torch.from_numpy(T).to(dev), depth.clone(), colour.clone()))
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
tt, ti = [], []
for s in range(24):
    R, T, depth, colour = frames[s % 4]
    ev[0].record()
    be.set_transforms(R, T)
    ev[1].record()
    st = be.integrate(depth, colour, False, 1.0)
    ev[2].record()
    torch.cuda.synchronize()
    if s >= 4:
        tt.append(ev[0].elapsed_time(ev[1])); ti.append(ev[1].elapsed_time(ev[2]))
print(f"shard 0 of {N}: transforms {np.median(tt) * 1e3:.1f} us, cull+fused {np.median(ti) * 1e3:.1f} us, stats {st.tolist()}")
if timing:
    L = _capi.load()
    buf = (ctypes.c_ulonglong * 4096)()
    L.ofx_debug_cta(buf)
    a = np.array(list(buf), dtype=np.int64).reshape(1024, 4)
    a = a[a[:, 0] > 0]
    t0 = a[:, 0].min()
    print("CTAs", len(a), "kernel span us", (a[:, 2].max() - t0) / 1e3)
    print("entry  us: min %.1f med %.1f max %.1f" % tuple(np.percentile((a[:, 0] - t0) / 1e3, [0, 50, 100])))
    print("first data us after entry: min %.1f med %.1f max %.1f" % tuple(np.percentile((a[:, 1] - a[:, 0]) / 1e3, [0, 50, 100])))
    print("end    us: min %.1f p10 %.1f med %.1f p90 %.1f max %.1f" % tuple(np.percentile((a[:, 2] - t0) / 1e3, [0, 10, 50, 90, 100])))
    print("bricks per CTA: min %d med %d max %d" % tuple(np.percentile(a[:, 3], [0, 50, 100])))
    print("us per brick per CTA: min %.2f med %.2f max %.2f" % tuple(np.percentile((a[:, 2] - a[:, 1]) / 1e3 / np.maximum(a[:, 3], 1), [0, 50, 100])))
