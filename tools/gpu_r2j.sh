#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_host_gpu.py -m gpu -q > gpurun_out/r2j_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2j_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/r2j_tests.log | cut -c1-500 | head -30
