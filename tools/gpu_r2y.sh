#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --workload cfg5chrom --no-cpu-baseline --no-parity-check --warmup 2 --steps 2 > gpurun_out/r2y_cfg5.log 2> gpurun_out/r2y.err; echo "rc $?"; tail -n 3 gpurun_out/r2y.err
python tools/show2.py gpurun_out/r2y_cfg5.log
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -q 2>&1 | tail -2
