#!/bin/bash
# the whole GPU suite and the default workload's bench line without the extra configs / CPU leg
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -x --timeout 150 > gpurun_out/all_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/all_tests.log
grep -n "^FAILED\|^ERROR\|passed\|failed\|tests rc\|^E  " gpurun_out/all_tests.log | cut -c1-300 | head -20
timeout 300 python bench.py --no-extra-configs --no-cpu-baseline > gpurun_out/short_bench.log 2> gpurun_out/short_bench.err; echo "bench rc $?"
tail -c 500 gpurun_out/short_bench.err
python tools/show2.py gpurun_out/short_bench.log
