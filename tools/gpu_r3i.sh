#!/bin/bash
# Launch list of a GoogLeNet step
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r3i_launches_googlenet.csv \
  python tools/model_rates.py googlenet > gpurun_out/r2x_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r2x_ncu.log
python - <<'PY'
import csv,collections,re
lines=[l for l in open('gpurun_out/r3i_launches_googlenet.csv') if l.startswith('"')]
agg=collections.defaultdict(lambda:[0,0.0]); tot=0
for row in csv.DictReader(lines):
    if row.get('Metric Name')!='gpu__time_duration.sum': continue
    name=re.sub(r'\(.*','',row['Kernel Name'])[:60]
    v=float(row['Metric Value'].replace(',','')); u=row['Metric Unit']
    v = v/1000 if u=='ns' else (v*1000 if u=='ms' else v)
    agg[name][0]+=1; agg[name][1]+=v; tot+=v
for k,(c,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:22]:
    print(f"{t:10.1f} us {c:5d} {t/c:8.1f} avg {t/tot*100:5.1f}%  {k}")
print(tot)
PY
echo ALLDONE
