#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.log 2>&1; rc=$?
tail -8 gpurun_out/r2_gpu_tests.log; echo "tests rc=$rc"
ALSO="c2,c1" timeout 600 bash tools/ab_bench.sh base g1s1c4
for wl in c1 c2 c2-dense; do timeout 300 python tools/timing_probe.py $wl 2>gpurun_out/tp_$wl.err | grep -v '^{' ; done
