#!/bin/bash
# round 2, call f: GPU tests with the integral-free solver steps and the tie-tolerant minima; bench with extra configs
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2f_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2f_tests.log
tail -12 gpurun_out/r2f_tests.log | cut -c1-200
timeout 600 python bench.py --no-cpu-baseline --warmup 3 --steps 3 > gpurun_out/r2f_bench.log 2> gpurun_out/r2f_bench.err; echo "rc $?"
tail -c 600 gpurun_out/r2f_bench.err
python tools/show2.py gpurun_out/r2f_bench.log
