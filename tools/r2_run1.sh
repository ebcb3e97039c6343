#!/bin/bash
# round 2, GPU call 1: parity tests, A/B of the pipeline shapes, ncu of the new fused kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.log 2>&1; rc=$?
tail -15 gpurun_out/r2_gpu_tests.log; echo "tests rc=$rc"
ALSO="c2,c1" timeout 1200 bash tools/ab_bench.sh base g1s1c4 g1s2c3 g3s4c1 g3s5c1 nodepth
if [ $rc -eq 0 ]; then
  timeout 300 python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --also "" > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_warped_persistent -s 4 -c 2 -f -o gpurun_out/r2_fused python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --also "" > gpurun_out/ncu1.log 2>&1
  echo "ncu rc=$?"; tail -3 gpurun_out/ncu1.log
fi
