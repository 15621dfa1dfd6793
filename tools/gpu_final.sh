#!/bin/bash
# end of the round: the whole GPU suite, smoke(), the driver's default bench command, the reference arm, and the launch list
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q --timeout 150 > gpurun_out/final_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/final_tests.log
grep -n "^FAILED\|^ERROR\|passed\|failed\|tests rc" gpurun_out/final_tests.log | cut -c1-300 | head -20
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/final_bench.log 2> gpurun_out/final_bench.err; echo "bench rc $?"
python tools/show2.py gpurun_out/final_bench.log
timeout 400 python bench.py --impl reference > gpurun_out/final_ref.log 2> gpurun_out/final_ref.err; echo "ref rc $?"; cut -c1-300 gpurun_out/final_ref.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 200 -c 80 --csv --log-file gpurun_out/final_launches.csv python bench.py --no-cpu-baseline --no-extra-configs --no-parity-check --warmup 4 --steps 1 > gpurun_out/final_ncu.log 2>&1; echo "ncu rc $?"
