#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_preprocess_gpu.py tests/test_host_gpu.py -m gpu -q > gpurun_out/r2v_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2v_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/r2v_tests.log | cut -c1-400 | head -20
