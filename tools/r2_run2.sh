#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.log 2>&1; rc=$?
tail -8 gpurun_out/r2_gpu_tests.log; echo "tests rc=$rc"
for wl in c1 c2 c2-dense; do
  OFX_LIB=$PWD/occlusionfusion_b200/lib/libofx_g1s1c4.so timeout 300 python bench.py --workload $wl --steps 3 --warmup 3 --no-graph --no-cpu-baseline --also "" > gpurun_out/plain_$wl.log 2>&1 &&
  OFX_LIB=$PWD/occlusionfusion_b200/lib/libofx_g1s1c4.so timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_$wl.csv python bench.py --workload $wl --steps 3 --warmup 3 --no-graph --no-cpu-baseline --also "" > gpurun_out/ncu_l_$wl.log 2>&1
  echo "launch list $wl rc=$?"
done
OFX_LIB=$PWD/occlusionfusion_b200/lib/libofx_g1s1c4.so timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_warped_persistent -s 4 -c 2 -f -o gpurun_out/r2_fused_g1s1c4 python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --also "" > gpurun_out/ncu1.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu1.log
