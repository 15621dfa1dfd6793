#!/bin/bash
# round 2, call c: EM-loop tests with the new goldens, DMMA probe, solver iteration histogram in the evolved state, ncu captures
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_em_loop_gpu.py -m gpu -x -q -s > gpurun_out/r2c_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2c_tests.log
./tools/micro/dmma_probe > gpurun_out/r2c_dmma.txt 2>&1
CNF2_ENGINE_LIB=$PWD/cnf2freq_b200/lib/libcnf2_engine_hist.so timeout 300 python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 4 --steps 1 > gpurun_out/r2c_hist.log 2> gpurun_out/r2c_hist.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mstep_infprobs --launch-skip 45 --launch-count 1 -o gpurun_out/r2c_infprobs python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 5 --steps 1 > gpurun_out/r2c_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:warp_estep --launch-skip 25 --launch-count 1 -o gpurun_out/r2c_warp python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 3 --steps 1 > gpurun_out/r2c_ncu2.log 2>&1
tail -3 gpurun_out/r2c_tests.log; cat gpurun_out/r2c_dmma.txt; tail -c 1500 gpurun_out/r2c_hist.err
