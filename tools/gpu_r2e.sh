#!/bin/bash
# round 2, call e: all GPU tests, ablations of the warp kernel (E-step only), ncu capture of the new warp kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2e_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2e_tests.log
tail -15 gpurun_out/r2e_tests.log
B="python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 1 --steps 2"
for V in noturns noturns_notransfer nonsatomic; do
  CNF2_ENGINE_LIB=$PWD/cnf2freq_b200/lib/libcnf2_engine_$V.so timeout 300 $B > gpurun_out/r2e_abl_$V.log 2>> gpurun_out/r2e_abl.err
done
python tools/show2.py gpurun_out/r2e_abl_*.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:warp_estep --launch-skip 12 --launch-count 1 -o gpurun_out/r2e_warp python bench.py --no-parity-check --no-extra-configs --no-cpu-baseline --warmup 1 --steps 1 > gpurun_out/r2e_ncu.log 2>&1
