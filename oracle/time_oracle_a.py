"""TEST INFRASTRUCTURE ONLY -- times "Oracle-A" of SURVEY.md section 6: the reference's OWN Python (tsdf.py CPU branch +
warpfield.py, imported in place from /root/reference through oracle/ref_shims.py) on BASELINE.json's configs[0]
(128^3 TSDF, 640x480 synthetic depth, ~400-node graph, K = 4).  Runs in the build container only (the reference does not
travel to the GPU box); writes profiles/r2_oracle_a_build_container.json.

    python oracle/time_oracle_a.py [n_warped_frames]

Drives the reference objects as fusion.py:57-83,176-179 does: TSDFVolume, integrate(source frame), WarpField,
update_transformations, integrate(warped frame).  WarpField.deform is hard-wired to optimizer.deform_ED (lietorch, missing
submodule); the stand-in of oracle/make_golden.py evaluates the blend with the reference's own numba deform_lbs."""
import json
import os
import sys
import tempfile
import time
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import ref_shims                                        # noqa: E402
from make_golden import LbsOptimizer                    # noqa: E402
from occlusionfusion_b200.synthetic import Workload    # noqa: E402


def main():
    n_warped = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    ref_tsdf, ref_wf = ref_shims.import_reference()
    w = Workload("c0")
    tmp = tempfile.mkdtemp(prefix="ofx_oracle_a_")
    fopt = types.SimpleNamespace(voxel_dim=[int(d) for d in w.dims], source_frame=0, skip_rate=1, datadir=tmp, gpu=False)
    t0 = time.perf_counter()
    vol = ref_tsdf.TSDFVolume(w.bbox, w.max_depth, w.intr, fopt, None)
    t_ctor = time.perf_counter() - t0
    assert tuple(int(d) for d in vol._vol_dim) == tuple(int(d) for d in w.dims), (vol._vol_dim, w.dims)
    t0 = time.perf_counter()
    vol.integrate({"im": w.frame(0), "id": 0})
    t_source = time.perf_counter() - t0
    graph = types.SimpleNamespace(nodes=w.nodes, graph_generation_parameters={"node_coverage": w.coverage})
    wf = ref_wf.WarpField(graph, vol, None)
    wf.graph_neighbours = w.K
    wf.optimizer = LbsOptimizer(wf, ref_wf.WarpField)
    vol.warpfield = wf
    times, updates = [], []
    for f in range(1, n_warped + 1):
        est = w.estimate(f)
        im = w.frame(f)
        w_before = vol._weight_vol_cpu.sum(dtype=np.float64)
        t0 = time.perf_counter()
        wf.update_transformations(est)
        vol.integrate({"im": im, "id": f})
        times.append(time.perf_counter() - t0)
        updates.append(int(round(vol._weight_vol_cpu.sum(dtype=np.float64) - w_before)))      # obs_weight = 1 per update
    steady = times[1:] if len(times) > 1 else times
    upd = updates[1:] if len(updates) > 1 else updates
    out = {
        "what": "Oracle-A: the reference's own Python (tsdf.py CPU branch + warpfield.py) imported in place, BASELINE configs[0]",
        "workload": f"c0: {w.dims[0]}x{w.dims[1]}x{w.dims[2]} voxels, {w.H}x{w.W} depth, {w.M} nodes, K={w.K}",
        "host": {"cores": os.cpu_count(), "where": "build container (no GPU)"},
        "constructor_s": t_ctor, "source_frame_s": t_source,
        "first_warped_frame_s": times[0], "first_warped_frame_includes": "KD-tree skinning of every voxel (cached afterwards) + numba JIT",
        "steady_warped_frame_s": float(np.median(steady)), "steady_frames": len(steady),
        "updates_per_frame": float(np.mean(upd)), "updates_per_s_steady": float(np.mean(upd) / np.median(steady)),
        "stand_ins": "pykdtree -> scipy cKDTree, deform_ED -> the reference's numba deform_lbs (oracle/ref_shims.py, oracle/make_golden.py)",
    }
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", "r2_oracle_a_build_container.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
