#!/bin/bash
# quick experiment round on the GPU box: M-step + engine parity tests, then one bench line
mkdir -p gpurun_out
python -m pytest tests/test_mstep_gpu.py tests/test_engine_gpu.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/exp_tests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/exp_bench.log 2> gpurun_out/exp_bench.err
tail -c 3000 gpurun_out/exp_bench.err
python tools/show.py gpurun_out/exp_bench.log
