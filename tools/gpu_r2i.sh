#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_host_gpu.py -m gpu -q -x > gpurun_out/r2i_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2i_tests.log
grep -n "^FAILED\|^E  \|passed\|failed" gpurun_out/r2i_tests.log | cut -c1-400 | head -30
cd tests/golden/host/refdata && ../../../../host/cnf2freq_gpu --mapfile demoplantimpute.map --pedfile demoplantimpute.ped --genfile demoplantimpute.gen --count 3 --output /tmp/demo.out > /root/repo/gpurun_out/r2i_demo.stdout 2> /root/repo/gpurun_out/r2i_demo.stderr; echo "demo rc $?"; cp /tmp/demo.out /root/repo/gpurun_out/r2i_demo.out
grep -v "^Creating" /root/repo/gpurun_out/r2i_demo.stdout | head -50; tail -5 /root/repo/gpurun_out/r2i_demo.stderr
