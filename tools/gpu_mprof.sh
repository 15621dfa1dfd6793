#!/bin/bash
# ncu evidence for the M-step kernels: launch list of one EM iteration, full capture of one launch of each solver kernel
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --chroms 3"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:mstep --csv --log-file gpurun_out/m_launches.csv $B > gpurun_out/m_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mstep_infprobs -c 1 -o gpurun_out/m_infprobs $B > gpurun_out/m_i.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mstep_haploweight -s 1 -c 1 -o gpurun_out/m_hw $B > gpurun_out/m_h.log 2>&1
ls -la gpurun_out | tail -5
