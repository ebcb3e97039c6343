#!/bin/bash
# A/B timing of experimental builds of libofx.so (run on the GPU box): tools/ab_bench.sh name1 name2 ...
# Variants are built beforehand (CPU box) with
#   python -m occlusionfusion_b200.build --out=occlusionfusion_b200/lib/libofx_<name>.so -DOFX_...
for v in "$@"; do
  lib=occlusionfusion_b200/lib/libofx_$v.so
  [ "$v" = base ] && lib=occlusionfusion_b200/lib/libofx.so
  OFX_LIB=$PWD/$lib python bench.py --steps 30 --warmup 5 --no-cpu-baseline --also "${ALSO-}" > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/ab_{v}.json"))
    extra = {k: round(x["kernel_ms"], 4) for k, x in d.get("also", {}).items()}
    print(f"{v:24s} ms/step {d['ms_per_step']:.4f} kernel {d['roofline']['kernel_ms']:.4f} e2e {d['e2e']['ms_per_step']:.3f} upd {d['updated_per_frame']:.0f} {extra}")
except Exception as e:
    print(v, "FAILED", e, open(f"gpurun_out/ab_{v}.err").read()[-400:])
PY
done
