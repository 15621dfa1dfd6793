#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2h_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2h_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/r2h_tests.log | cut -c1-220 | head -30
timeout 600 python bench.py --no-cpu-baseline --no-extra-configs --warmup 3 --steps 3 > gpurun_out/r2h_bench.log 2> gpurun_out/r2h_bench.err; echo "rc $?"
python tools/show2.py gpurun_out/r2h_bench.log
