timeout 600 python -m pytest tests/test_engine_multigpu.py -m gpu -x -q > gpurun_out/mg_tests.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/mg_bench2.log 2>&1
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/ref_arm.log 2>&1
