#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
N=8
for v in base pdlsh base pdlsh; do
  lib=occlusionfusion_b200/lib/libofx.so; [ $v = pdlsh ] && lib=occlusionfusion_b200/lib/libofx_pdlsh.so
  OFX_LIB=$PWD/$lib timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 --also "" --no-e2e --no-parity > gpurun_out/r2_n8_$v.json 2> gpurun_out/r2_n8_$v.err
  python -c "
import json;d=json.load(open('gpurun_out/r2_n8_$v.json'));print('$v', d['ms_per_step'], d['roofline']['kernel_ms'])"
done
