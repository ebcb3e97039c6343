#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.log 2>&1; rc=$?
tail -4 gpurun_out/r2_gpu_tests.log; echo "tests rc=$rc"
ALSO="c2,c1" timeout 600 bash tools/ab_bench.sh base
if [ $rc -eq 0 ]; then
  timeout 300 python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --no-e2e --also "" > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_warped_persistent -s 4 -c 1 -f -o gpurun_out/r2_fused_v12b python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --no-e2e --also "" > gpurun_out/ncu1.log 2>&1
  echo "ncu rc=$?"
fi
