timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/exp_tests.log 2>&1
