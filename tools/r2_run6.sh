#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.log 2>&1; rc=$?
tail -6 gpurun_out/r2_gpu_tests.log; echo "tests rc=$rc"
timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench rc=$?"
tail -5 gpurun_out/r2_bench_n1.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n1.json'))
print('ms',d['ms_per_step'],'frac',d['roofline']['frac'],'e2e',d['e2e']['ms_per_step'],'clocks',d['clocks'],'parity',d['parity'], 'fs', d['free_space_share_of_observed_voxels'])
for k,v in d['also'].items():
    if k=='c3': print(k, {x:v[x] for x in ['ms_per_frame_all_in','frame_loop_ms_per_frame','update_graph_ms_median','nodes']})
    else: print(k, v['ms_per_step'], v['kernel_ms'], 'e2e', v['e2e']['ms_per_step'], 'frac', v['roofline']['frac'], 'parity', v['parity'], 'fs', v['free_space_share_of_observed_voxels'])
print(d['cpu_baseline']['value'], d['cpu_baseline']['cores'])
PY
