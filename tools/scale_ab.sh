#!/bin/bash
# tools/scale_ab.sh N tag [ENV=VALUE ...]: one pipelined bench run on N GPUs (run on the GPU box); BENCH_ARGS='--workload c4' adds bench.py arguments
N=$1; tag=$2; shift 2
env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
  bench.py --gpus $N --steps 50 --warmup 5 ${BENCH_ARGS-} > gpurun_out/bench_${tag}_n$N.json 2> gpurun_out/bench_${tag}_n$N.err
python - "$N" "$tag" <<'PY'
import json, sys
n, tag = sys.argv[1:3]
try:
    d = json.loads(open(f"gpurun_out/bench_{tag}_n{n}.json").read().strip().splitlines()[-1])
    print(f"n{n} {tag:12s} value {d['value']:.4g} ms/step {d['ms_per_step']:.4f} kernel {d['roofline']['kernel_ms']:.4f} e2e {(d.get('e2e') or {}).get('ms_per_step', float('nan')):.3f}")
except Exception as e:
    print(n, tag, "FAILED", e, open(f"gpurun_out/bench_{tag}_n{n}.err").read()[-600:])
PY
