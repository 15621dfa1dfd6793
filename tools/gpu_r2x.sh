#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:warp_estep --launch-skip 4 --launch-count 1 -o gpurun_out/r2x_fullsib python bench.py --workload fullsib --no-parity-check --no-cpu-baseline --warmup 2 --steps 1 > gpurun_out/r2x_ncu1.log 2>&1
timeout 900 python bench.py > gpurun_out/r2x_bench.log 2> gpurun_out/r2x_bench.err; echo "rc $?"
python tools/show2.py gpurun_out/r2x_bench.log
