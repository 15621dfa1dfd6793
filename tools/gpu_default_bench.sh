#!/bin/bash
# the driver's default command at the round's last commit
mkdir -p gpurun_out
timeout 400 python bench.py > gpurun_out/final_bench.log 2> gpurun_out/final_bench.err; echo "bench rc $?"
python tools/show2.py gpurun_out/final_bench.log
