#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_engine_gpu.py -m gpu -q -k "reference_golden" > gpurun_out/r2u_tests.log 2>&1; tail -3 gpurun_out/r2u_tests.log | cut -c1-300
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2u_memcheck.log 2>&1; echo "memcheck rc $?"; tail -4 gpurun_out/r2u_memcheck.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_em_loop_gpu.py -m gpu -q -k "free_running and mstep_inv_fullsib" > gpurun_out/r2u_memcheck2.log 2>&1; echo "memcheck2 rc $?"; tail -4 gpurun_out/r2u_memcheck2.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2u_racecheck.log 2>&1; echo "racecheck rc $?"; tail -4 gpurun_out/r2u_racecheck.log
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 ./host/cnf2freq_gpu --mapfile tests/golden/host/f2.map --pedfile tests/golden/host/f2.ped --genfile tests/golden/host/f2.gen --count 2 --output /tmp/o.txt > gpurun_out/r2u_memcheck3.log 2>&1; echo "memcheck3 rc $?"; tail -3 gpurun_out/r2u_memcheck3.log
