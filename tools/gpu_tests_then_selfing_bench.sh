#!/bin/bash
# whole GPU suite (pytest-timeout per test), then the SELFING workload's bench line
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -x --timeout 120 > gpurun_out/all_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/all_tests.log
grep -n "^FAILED\|^ERROR\|passed\|failed\|tests rc\|^E  " gpurun_out/all_tests.log | cut -c1-300 | head -20
bash tools/gpu_selfing_bench.sh
