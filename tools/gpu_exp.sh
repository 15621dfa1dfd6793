timeout 600 python -m pytest tests/test_mstep_gpu.py -m gpu -x -q > gpurun_out/exp_ms.log 2>&1
