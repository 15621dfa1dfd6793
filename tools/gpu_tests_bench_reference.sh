#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/final_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/final_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/final_tests.log | cut -c1-300 | head -20
timeout 900 python bench.py > gpurun_out/final_bench.log 2> gpurun_out/final_bench.err; echo "rc $?"
python tools/show2.py gpurun_out/final_bench.log
timeout 600 python bench.py --impl reference > gpurun_out/final_ref.log 2> gpurun_out/final_ref.err; echo "ref rc $?"; cut -c1-400 gpurun_out/final_ref.log
