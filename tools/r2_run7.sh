#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
nvidia-smi -L | head -3
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench N=$N rc=$?"
tail -8 gpurun_out/r2_bench_n$N.err
python - $N <<'PY'
import json,sys
N=sys.argv[1]
d=json.load(open(f'gpurun_out/r2_bench_n{N}.json'))
print('ms',d['ms_per_step'],'value',d['value'],'frac',d['roofline']['frac'],'e2e',d['e2e'],'clocks',d['clocks'],'parity',{k:d['parity'][k] for k in ['checked','mismatches']})
for k,v in d['also'].items():
    print(k, v['ms_per_step'], v['kernel_ms'], 'e2e', v['e2e']['ms_per_step'], 'frac', v['roofline']['frac'], 'parity', v['parity'])
PY
