#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 200 -c 80 --csv --log-file gpurun_out/r2t_launches.csv python bench.py --no-cpu-baseline --no-extra-configs --no-parity-check --warmup 4 --steps 1 > gpurun_out/r2t_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/r2t_launches.csv')) if len(r)>10 and r[0].isdigit()]
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows:
    name=r[4].split('(')[0][-40:]; v=float(r[-1].replace(',',''))
    unit=r[-2]
    if unit=='usecond': v/=1e3
    elif unit=='nsecond': v/=1e6
    elif unit=='second': v*=1e3
    agg[name][0]+=1; agg[name][1]+=v
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1]): print("%-42s n=%3d total %8.2f ms avg %7.3f ms"%(k,n,t,t/n))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mstep_infprobs --launch-skip 45 --launch-count 1 -o gpurun_out/r2t_infprobs python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 5 --steps 1 > gpurun_out/r2t_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mstep_haploweight --launch-skip 48 --launch-count 1 -o gpurun_out/r2t_haplo python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 5 --steps 1 > gpurun_out/r2t_ncu2.log 2>&1
