#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for n in 8 4; do timeout 300 python tools/shard_probe.py $n timing; done
timeout 300 python tools/shard_probe.py 8 > /dev/null 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_shard8.csv python tools/shard_probe.py 8 > gpurun_out/ncu_shard.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r2_launches_shard8.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if r[0]=='ID'][0]
h=rows[hdr]; iK=h.index('Kernel Name'); iV=h.index('Metric Value')
seq=[(r[iK].split('(')[0][-50:], float(r[iV].replace(',',''))) for r in rows[hdr+1:]]
for k,v in seq[-8:]: print(f"   {k:50s} {v/1000:8.2f} us")
PY
