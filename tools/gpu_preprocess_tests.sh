#!/bin/bash
# row f3 after the factorised ranking: pre-processing tests and the host program's end-to-end runs
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_preprocess_gpu.py tests/test_host_gpu.py -m gpu -q -s --timeout 120 > gpurun_out/preprocess_tests.log 2>&1; echo "rc $?" >> gpurun_out/preprocess_tests.log
grep -n "literal sums walked\|passed\|failed\|^FAILED\|^E  \|rc " gpurun_out/preprocess_tests.log | cut -c1-300 | head -30
