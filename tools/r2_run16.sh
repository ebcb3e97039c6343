#!/bin/bash
# final multi-GPU record of the round: tools/r2_run16.sh N  (the driver's own command line)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
  bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r2_final_bench_n$N.json 2> gpurun_out/r2_final_bench_n$N.err
echo "rc=$?"
python - "$N" <<'PY'
import json, sys
n = sys.argv[1]
d = json.loads(open(f"gpurun_out/r2_final_bench_n{n}.json").read().strip().splitlines()[-1])
print(f"n{n} value {d['value']:.4g} ms/step {d['ms_per_step']:.4f} e2e {d['e2e']['ms_per_step']:.3f} parity {d.get('parity')} clocks {d.get('clocks')}")
for k, v in (d.get("also") or {}).items():
    print("  also", k, {x: v.get(x) for x in ("ms_per_step", "value", "parity", "roofline_frac")})
PY
