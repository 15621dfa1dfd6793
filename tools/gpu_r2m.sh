#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2m_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2m_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/r2m_tests.log | cut -c1-300 | head -20
timeout 600 python bench.py --no-cpu-baseline --warmup 3 --steps 3 > gpurun_out/r2m_bench.log 2> gpurun_out/r2m_bench.err; echo "rc $?"
python tools/show2.py gpurun_out/r2m_bench.log
