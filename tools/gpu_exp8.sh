nvidia-smi -L | wc -l > gpurun_out/mg8_gpus.log 2>&1
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/mg_bench$n.log 2>&1
done
