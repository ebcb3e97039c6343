#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 300 python -m pytest tests/test_z_graph_edges_gpu.py tests/test_z_graph_gpu.py tests/test_z_mc_gpu.py -x -q -m gpu 2>&1 | tail -8
timeout 300 python tools/update_graph_probe.py 2>&1 | tail -18
timeout 300 python - <<'PY'
import cProfile, pstats, os, sys, io
sys.argv=['x']
src=open('tools/update_graph_probe.py').read().split("for name in (\"get_canonical_model\"")[0]
exec(src)
frames_=frames
for f in range(1, 10):
    est = w.estimate(f, wf.graph.nodes); wf.update_transformations(est); vol.integrate({"im": frames_[f % 8], "id": f})
torch.cuda.synchronize()
pr=cProfile.Profile(); pr.enable()
vol.clear(); wf.update_graph()
torch.cuda.synchronize()
pr.disable()
s=io.StringIO(); pstats.Stats(pr,stream=s).sort_stats('cumulative').print_stats(45); print(s.getvalue()[:7000])
PY
