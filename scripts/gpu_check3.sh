#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python scripts/bench_ops.py --steps 3 --warmup 2 > gpurun_out/ops2.log 2>&1; grep "^{" gpurun_out/ops2.log | cut -c1-700
python scripts/bench_ops.py --steps 1 --warmup 1 --shuffle-rows 1000 > gpurun_out/plain_ops.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_ops.csv python scripts/bench_ops.py --steps 1 --warmup 1 --shuffle-rows 1000 > gpurun_out/ncu_ops.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(l for l in open("gpurun_out/launches_ops.csv") if not l.startswith("=="))]
hdr=rows[0]
ik=hdr.index("Kernel Name"); iv=hdr.index("Metric Value")
agg={}
order=[]
for r in rows[1:]:
    if len(r)<=iv: continue
    k=r[ik].split("(")[0][-40:]
    try: v=float(r[iv].replace(",",""))
    except: continue
    if k not in agg: order.append(k)
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=v
for k in order: print(f"{k:42s} n={agg[k][0]:3d} total_ms={agg[k][1]/1e6:9.3f}")
PY
