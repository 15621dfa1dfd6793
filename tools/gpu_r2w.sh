#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2w_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2w_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/r2w_tests.log | cut -c1-300 | head -30
timeout 600 python bench.py --workload fullsib --no-cpu-baseline --no-parity-check --warmup 2 --steps 3 > gpurun_out/r2w_fullsib.log 2> gpurun_out/r2w.err; echo "rc $?"
python tools/show2.py gpurun_out/r2w_fullsib.log
