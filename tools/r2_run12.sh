#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 300 python tools/update_graph_probe.py > /dev/null 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_update_graph.csv python tools/update_graph_probe.py > gpurun_out/ncu_ug.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/r2_launches_update_graph.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if r[0]=='ID'][0]
h=rows[hdr]; iK=h.index('Kernel Name'); iV=h.index('Metric Value')
seq=[(r[iK].split('(')[0][-48:], float(r[iV].replace(',',''))) for r in rows[hdr+1:]]
# find last update_graph: from the last 'truncated_region' kernel to the next fused kernel
idx=[i for i,(k,v) in enumerate(seq) if 'truncated_region' in k]
start=idx[-2] if len(idx)>=2 else idx[-1]
end=next(i for i in range(start, len(seq)) if 'fused_warped' in seq[i][0])
tot=collections.OrderedDict()
for k,v in seq[start:end]:
    tot.setdefault(k,[0,0.0]); tot[k][0]+=1; tot[k][1]+=v
s=0
for k,(n,v) in tot.items():
    print(f"  {k:50s} x{n:3d} {v/1000:9.1f} us"); s+=v
print("  device total", s/1000, "us over", end-start, "launches")
PY
