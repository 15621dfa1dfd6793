CNF2_DEBUG=1 timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/exp_a.log 2>&1
