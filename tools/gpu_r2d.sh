#!/bin/bash
# round 2, call d: all GPU tests with the segmented warp kernel + fused turn scoring, bench (main workload), segment experiments
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2d_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2d_tests.log
tail -15 gpurun_out/r2d_tests.log
B="python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 2 --steps 2"
timeout 300 $B > gpurun_out/r2d_bench.log 2> gpurun_out/r2d_bench.err; echo "rc $?"
for S in 8 16 32 64; do CNF2_SEG=$S timeout 300 $B > gpurun_out/r2d_bench_seg$S.log 2>> gpurun_out/r2d_bench.err; done
python tools/show2.py gpurun_out/r2d_bench.log gpurun_out/r2d_bench_seg*.log
