#!/bin/bash
# The SELFING workload alone (literal kernel, 256 states): how long one EM iteration takes at the size `configs` uses.
mkdir -p gpurun_out
timeout 240 python bench.py --workload selfing --no-cpu-baseline --no-parity-check --warmup 2 --steps 2 > gpurun_out/selfing_bench.json 2> gpurun_out/selfing_bench.err; echo "rc $?"
tail -3 gpurun_out/selfing_bench.err | cut -c1-400
python tools/show2.py gpurun_out/selfing_bench.json
