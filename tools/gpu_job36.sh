#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ncu_stats_target.py > gpurun_out/ncu_stats_plain.log 2>&1 || { echo plain failed; tail -3 gpurun_out/ncu_stats_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_stats.csv python tools/ncu_stats_target.py > gpurun_out/ncu_stats.log 2>&1
echo "ncu exit=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open("gpurun_out/launches_stats.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); gi=h.index("Grid Size")
for r in rows[1:][-18:]: print(r[ki][:60], r[gi], r[vi])
PY
