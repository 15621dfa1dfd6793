#!/bin/bash
# Row g1 on the GPU: the SELFING tests first (their own log), then the whole GPU suite.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_selfing_gpu.py -m gpu -q -x > gpurun_out/selfing_tests.log 2>&1; echo "selfing rc $?" >> gpurun_out/selfing_tests.log
tail -40 gpurun_out/selfing_tests.log | cut -c1-400
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/all_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/all_tests.log
grep -n "^FAILED\|^ERROR\|passed\|failed\|tests rc" gpurun_out/all_tests.log | cut -c1-300 | head -30
