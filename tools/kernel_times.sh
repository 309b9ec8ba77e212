#!/bin/bash
# per-kernel durations of one warm decode step (single-stream profile decoder) -> gpurun_out/<tag>_ktimes.csv
tag=${1:-kt}
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:'k_' --launch-skip ${NCU_SKIP:-33} --launch-count ${NCU_COUNT:-11} --csv \
  --log-file gpurun_out/${tag}_ktimes.csv python tools/quick_bench.py 64 8 "${2:-:fancy_upsampling=1 :profile=1}" > gpurun_out/${tag}_ktimes.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(l for l in open("gpurun_out/${tag}_ktimes.csv") if l.startswith('"'))]
h=rows[0]; ik=h.index("Kernel Name"); im=h.index("Metric Name"); iv=h.index("Metric Value"); iid=h.index("ID")
d={}
for r in rows[1:]:
    d.setdefault((r[iid], r[ik].split('(')[0]), {})[r[im]]=r[iv]
for (i,k),m in d.items():
    print(k.ljt(20) if False else k.ljust(20), " ".join("%s=%s"%(a.split('.')[0][-14:],b) for a,b in m.items()))
PY
