#!/bin/bash
mkdir -p gpurun_out
for N in 4 6 8 9; do
  CNF2_CTAS_PER_SM=$N timeout 300 python bench.py --workload fullsib --no-parity-check --no-cpu-baseline --warmup 2 --steps 2 > gpurun_out/r2s_fs$N.log 2>> gpurun_out/r2s.err
done
for N in 8 10 12; do
  CNF2_CTAS_PER_SM=$N timeout 300 python bench.py --no-extra-configs --no-parity-check --no-cpu-baseline --warmup 2 --steps 2 > gpurun_out/r2s_f2_$N.log 2>> gpurun_out/r2s.err
done
python tools/show2.py gpurun_out/r2s_fs*.log gpurun_out/r2s_f2_*.log
