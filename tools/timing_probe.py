"""Investigation only: runs bench.py's c2-dense frames with the OFX_TIMING build and prints where the
persistent kernel's warps wait.  tools/timing_probe.py (on the GPU box, after building libofx_timing.so)."""
import ctypes, os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "occlusionfusion_b200", "lib", "libofx_timing.so")
os.environ["OFX_LIB"] = lib
sys.path.insert(0, ROOT)
sys.argv = ["bench.py", "--steps", "10", "--warmup", "3", "--no-cpu-baseline", "--also", "", "--no-graph"]
import bench
bench.main()
from occlusionfusion_b200 import _capi
L = _capi.load()
out = (ctypes.c_ulonglong * 8)()
L.ofx_debug_timing(out, 0)
v = list(out)
print("producer: empty-wait ns total %d, count %d, avg %.0f ns" % (v[0], v[1], v[0] / max(v[1], 1)))
print("consumer: data-wait ns total %d, count %d, avg %.0f ns" % (v[2], v[3], v[2] / max(v[3], 1)))
print("consumer: brick work ns total %d, avg %.0f ns" % (v[4], v[4] / max(v[3], 1)))
print("producer: TMA issue->landed ns total %d, count %d, avg %.0f ns" % (v[5], v[6], v[5] / max(v[6], 1)))

import numpy as np
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
