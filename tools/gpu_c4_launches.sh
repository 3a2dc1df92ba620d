#!/bin/bash
# One gpurun call, ONE ncu invocation: the launch list (gpu__time_duration per launch) of the stock config-4 bench command.
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --no-cpu"
$B > gpurun_out/ncu_plain_c4.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c4.csv $B > gpurun_out/ncu_launches_c4.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(l for l in open("gpurun_out/launches_c4.csv") if l.startswith('"'))]
h=rows[0]; k=h.index("Kernel Name"); v=h.index("Metric Value")
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows[1:]:
    a=agg[r[k][:90]]; a[0]+=1; a[1]+=float(r[v].replace(",",""))
tot=sum(a[1] for a in agg.values())
for name,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:8]: print(f"{n:4d} launches {t/1e6:9.3f} ms {100*t/tot:5.1f}%  {name}")
PY
