#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-ll}
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -c 200 --csv \
    --log-file $O/${TAG}_launches.csv python bench.py --workload C2 --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(l for l in open("$O/${TAG}_launches.csv") if not l.startswith("=="))]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
d={}
for r in rows[1:]:
    if len(r)<=vi: continue
    d.setdefault((r[ii], r[ki].split("(")[0][-40:]),{})[r[mi]]=r[vi]
for (i,k),m in d.items():
    print(i,k, m.get("gpu__time_duration.sum"), m.get("smsp__inst_executed.sum"), m.get("smsp__issue_active.avg.pct_of_peak_sustained_active"))
PY
