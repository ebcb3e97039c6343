#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.log 2>&1; rc=$?
tail -4 gpurun_out/r2_gpu_tests.log; echo "tests rc=$rc"
timeout 900 python bench.py --steps 30 --warmup 5 --also c3 > gpurun_out/r2_bench_c3.json 2> gpurun_out/r2_bench_c3.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_c3.json'))
print('ms',d['ms_per_step'],'frac',d['roofline']['frac'],'e2e',d['e2e']['ms_per_step'])
v=d['also']['c3']; print({x:v[x] for x in ['ms_per_frame_all_in','frame_loop_ms_per_frame','update_graph_ms_median','update_graph_ms','nodes']})
PY
