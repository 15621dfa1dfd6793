#!/bin/bash
# round 2, call l: tests after the candidate index + 12 warps/SM; bench; launch list of one evolved EM iteration
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2l_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2l_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/r2l_tests.log | cut -c1-300 | head -20
timeout 600 python bench.py --no-cpu-baseline --no-extra-configs --warmup 3 --steps 3 > gpurun_out/r2l_bench.log 2> gpurun_out/r2l_bench.err; echo "rc $?"
python tools/show2.py gpurun_out/r2l_bench.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 200 -c 80 --csv --log-file gpurun_out/r2l_launches.csv python bench.py --no-cpu-baseline --no-extra-configs --no-parity-check --warmup 4 --steps 1 > gpurun_out/r2l_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/r2l_launches.csv')) if len(r)>10 and r[0].isdigit()]
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
