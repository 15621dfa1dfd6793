#!/bin/bash
# launch list (durations) of the M-step kernels of one short EM iteration
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:mstep --csv --log-file gpurun_out/m_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --chroms 3 > gpurun_out/m_l.log 2>&1
python - <<'PY'
import csv
for r in csv.reader(open('gpurun_out/m_launches.csv')):
    if len(r)>10 and r[0].isdigit() and int(r[0])<6: print(r[4][:40], r[-1])
PY
