timeout 600 python -m pytest tests/test_mstep_gpu.py -m gpu -x -q > gpurun_out/exp_ms.log 2>&1
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/exp_a.log 2>&1
