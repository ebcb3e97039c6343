"""Investigation only: runs one bench workload with the OFX_TIMING build (libofx_timing.so) and prints, per CTA of the LAST
fused_warped_persistent launch, where the time goes: python tools/timing_probe.py [workload]
(build first: python -m occlusionfusion_b200.build --out=occlusionfusion_b200/lib/libofx_timing.so -DOFX_TIMING=1 ...)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.environ.get("OFX_LIB") or os.path.join(ROOT, "occlusionfusion_b200", "lib", "libofx_timing.so")
os.environ["OFX_LIB"] = lib
wl = sys.argv[1] if len(sys.argv) > 1 else "c1"
sys.path.insert(0, ROOT)
sys.argv = ["bench.py", "--workload", wl, "--steps", "6", "--warmup", "3", "--no-cpu-baseline", "--also", "", "--no-graph", "--no-e2e"]
import bench
bench.main()
import numpy as np
from occlusionfusion_b200 import _capi
L = _capi.load()
buf = (ctypes.c_ulonglong * (2048 * 8))()
L.ofx_debug_cta(buf)
a = np.array(list(buf), dtype=np.int64).reshape(2048, 8)
a = a[a[:, 0] > 0]
t0 = a[:, 0].min()
pc = lambda x: "min %.1f med %.1f p90 %.1f max %.1f" % tuple(np.percentile(x, [0, 50, 90, 100]))
print(f"[{wl}] CTAs {len(a)}; kernel span {(a[:, 5].max() - t0) / 1e3:.1f} us (first entry -> last consumer exit)")
print("  entry (us after first entry):            ", pc((a[:, 0] - t0) / 1e3))
print("  wait for the cull  (entry -> past wait):  ", pc((a[:, 1] - a[:, 0]) / 1e3))
w = a[a[:, 2] > 0]
print("  first data landed (after the wait):      ", pc((w[:, 2] - w[:, 1]) / 1e3))
w = a[a[:, 3] > 0]
print(f"  CTAs with a brick: {len(w)}")
print("  first brick phase 1 (table + blend):     ", pc((w[:, 3] - w[:, 2]) / 1e3))
print("  first brick phases 2+3:                  ", pc((w[:, 4] - w[:, 3]) / 1e3))
print("  rest (more bricks, end marker):          ", pc((w[:, 5] - w[:, 4]) / 1e3))
print("  consumer exit (us after first entry):    ", pc((a[:, 5] - t0) / 1e3))
