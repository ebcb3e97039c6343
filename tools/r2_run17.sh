#!/bin/bash
# sweep of the launch-time row-table capacity of the persistent kernel (OFX_TAB_CAP override, investigation only)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
run() { # label, env...
  lab=$1; shift
  env "$@" OFX_DEBUG=1 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-parity --no-e2e --workload ${WL-c2-dense} --also "" > gpurun_out/sw.json 2> gpurun_out/sw.err
  python - "$lab" <<'PY'
import json, sys
d = json.load(open("gpurun_out/sw.json"))
dbg = [l for l in open("gpurun_out/sw.err") if l.startswith("[ofx]")]
print(f"{sys.argv[1]:28s} ms/step {d['ms_per_step']:.4f}   {dbg[-1].strip() if dbg else ''}")
PY
}
for wl in c2-dense c2; do
  export WL=$wl; echo "== $wl"
  run default
  for tc in 8 12 14 16 20 24 32; do run "tab_cap=$tc" OFX_TAB_CAP=$tc; done
done
