#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_engine_gpu.py -m gpu -q > gpurun_out/r2p_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2p_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/r2p_tests.log | cut -c1-300 | head -20
timeout 600 python bench.py --no-cpu-baseline --no-parity-check --warmup 3 --steps 3 > gpurun_out/r2p_bench.log 2> gpurun_out/r2p_bench.err; echo "rc $?"
python tools/show2.py gpurun_out/r2p_bench.log
