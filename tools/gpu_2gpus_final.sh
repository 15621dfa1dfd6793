#!/bin/bash
# 2 GPUs at the round's last state: the multi-GPU parity tests, the SELFING shard test, a 2-rank bench line
mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 400 python -m pytest tests/test_engine_multigpu.py tests/test_selfing_gpu.py -m gpu -q --timeout 150 > gpurun_out/final_tests_2gpu.log 2>&1; echo "tests rc $?" >> gpurun_out/final_tests_2gpu.log
tail -4 gpurun_out/final_tests_2gpu.log | cut -c1-300
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_2gpu.log 2> gpurun_out/final_bench_2gpu.err; echo "bench2 rc $?"
python tools/show2.py gpurun_out/final_bench_2gpu.log
