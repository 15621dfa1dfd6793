#!/bin/bash
# Row g1 on the GPU: the SELFING tests (E-step, EM iterations, pre-processing, host program), each under its own time limit.
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_selfing_gpu.py "tests/test_preprocess_gpu.py" "tests/test_host_gpu.py" -m gpu -q -k "selfing or True" > gpurun_out/selfing_tests.log 2>&1; echo "selfing rc $?" >> gpurun_out/selfing_tests.log
tail -40 gpurun_out/selfing_tests.log | cut -c1-900
