#!/bin/bash
# ncu captures of the fused kernel on three workloads (each after the same command ran plain) + launch lists
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for wl in c2-dense c2 c2-surface; do
  ARGS=(--workload $wl --steps 3 --warmup 3 --no-graph --no-cpu-baseline --no-e2e --no-parity --also "")
  timeout 300 python bench.py "${ARGS[@]}" > gpurun_out/plain_$wl.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_warped_persistent -s 4 -c 1 -f -o gpurun_out/r2_final_fused_$wl python bench.py "${ARGS[@]}" > gpurun_out/ncu_$wl.log 2>&1
  echo "$wl ncu rc=$?"
  timeout 300 python bench.py "${ARGS[@]}" > gpurun_out/plain2_$wl.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_final_launches_$wl.csv python bench.py "${ARGS[@]}" > gpurun_out/ncu_l_$wl.log 2>&1
  echo "$wl launches rc=$?"
done
