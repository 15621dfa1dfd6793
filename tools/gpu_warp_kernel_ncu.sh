#!/bin/bash
# ncu captures of the warp kernel on the full-sib shape and the final cfg3 kernel
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:warp_estep --launch-skip 4 --launch-count 1 -o gpurun_out/r2o_fullsib python bench.py --workload fullsib --no-parity-check --no-cpu-baseline --warmup 2 --steps 1 > gpurun_out/r2o_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:warp_estep --launch-skip 32 --launch-count 1 -o gpurun_out/r2o_cfg3 python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 3 --steps 1 > gpurun_out/r2o_ncu2.log 2>&1
ls -la gpurun_out/r2o*
