"""Investigation only: one GPU plays shard 0 of N (c2-dense by default) and reports the per-frame device time of the
per-frame launch sequence (frame preparation + node transforms, brick cull, fused kernel) and, with an OFX_TIMING build,
the per-CTA timeline of the fused kernel.
  tools/shard_probe.py N [name of an OFX_TIMING build: libofx_<name>.so]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
timing = len(sys.argv) > 2
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
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ti = []
for s in range(24):
    R, T, depth, colour = frames[s % 4]
    if os.environ.get("FLUSH", "1") == "1":
        flush.fill_(s & 0xFF)
    ev[0].record()
    st = be.integrate_warped(R, T, depth, colour, 1.0)
    ev[1].record()
    torch.cuda.synchronize()
    if s >= 4:
        ti.append(ev[0].elapsed_time(ev[1]))
print(f"shard 0 of {N} ({w.name}): prep+cull+fused {np.median(ti) * 1e3:.1f} us (min {np.min(ti) * 1e3:.1f}), stats {st.tolist()}")
# ---- K frames back to back: one stream (prep + cull + fused per frame) against the staged loop (ofx_stage_frame of frame
# t+1 on a second stream while frame t's fused kernel runs)
K = int(os.environ.get("LOOP", "200"))
lib = _capi.load()
def reset():
    _capi.check(lib.ofx_volume_reset(vol._h, _capi.stream_ptr(dev)))
    torch.cuda.synchronize()
def loop_plain():
    out = []
    for t in range(K):
        R, T, depth, colour = frames[t % 4]
        st = torch.zeros(4, dtype=torch.int64, device=dev)
        be.integrate_warped(R, T, depth, colour, 1.0, stats=st)
        out.append(st)
    return out
sB = torch.cuda.Stream(device=dev, priority=-1)
def loop_staged():
    sA = torch.cuda.current_stream(dev)
    ev_done = [torch.cuda.Event() for _ in range(2)]
    ev_staged = [torch.cuda.Event() for _ in range(2)]
    sB.wait_stream(sA)
    with torch.cuda.stream(sB):
        be.stage(*frames[0])
        ev_staged[0].record(sB)
    out = []
    for t in range(K):
        R, T, depth, colour = frames[t % 4]
        sA.wait_event(ev_staged[t % 2])
        st = torch.zeros(4, dtype=torch.int64, device=dev)
        be.integrate(depth, colour, False, 1.0, stats=st)
        ev_done[t % 2].record(sA)
        out.append(st)
        if t + 1 < K:
            with torch.cuda.stream(sB):
                if t >= 1:
                    sB.wait_event(ev_done[(t - 1) % 2])
                be.stage(*frames[(t + 1) % 4])
                ev_staged[(t + 1) % 2].record(sB)
    sA.wait_stream(sB)
    return out
res = {}
for name, fn in (("plain", loop_plain), ("staged", loop_staged), ("plain", loop_plain), ("staged", loop_staged)):
    reset()
    fn()                                        # warm
    reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    sts = fn()
    e1.record()
    torch.cuda.synchronize()
    tot = torch.stack(sts).cpu().numpy()
    vols = [a.copy() for a in vol.get_volume()]
    res.setdefault(name, []).append((e0.elapsed_time(e1) * 1e3 / K, tot, vols))
    print(f"  loop {name:7s}: {e0.elapsed_time(e1) * 1e3 / K:7.2f} us/frame over {K} frames")
a, b = res["plain"][0], res["staged"][0]
same = np.array_equal(a[1][:, [0, 3]], b[1][:, [0, 3]]) and all(np.array_equal(x, y) for x, y in zip(a[2], b[2]))
print("  staged loop == plain loop (updated / weight-sum counters of every frame, whole volume):", same)
if timing:
    L = _capi.load()
    buf = (ctypes.c_ulonglong * (2048 * 8))()
    L.ofx_debug_cta(buf)
    a = np.array(list(buf), dtype=np.int64).reshape(2048, 8)
    a = a[a[:, 0] > 0]
    t0 = a[:, 0].min()
    pc = lambda x: "min %.1f med %.1f p90 %.1f max %.1f" % tuple(np.percentile(x, [0, 50, 90, 100]))
    print(f"  CTAs {len(a)}; fused kernel span {(a[:, 5].max() - t0) / 1e3:.1f} us")
    print("  entry (us after first entry):        ", pc((a[:, 0] - t0) / 1e3))
    print("  first data landed (after the wait):  ", pc((a[:, 2] - a[:, 1]) / 1e3))
    print("  consumer exit (us after first entry):", pc((a[:, 5] - t0) / 1e3))
