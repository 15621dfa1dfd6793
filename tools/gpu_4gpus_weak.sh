#!/bin/bash
# 4 ranks, weak scaling of the bench workload at the round's last state
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_4gpu.log 2> gpurun_out/final_bench_4gpu.err; echo "bench4 rc $?"
python tools/show2.py gpurun_out/final_bench_4gpu.log
python - <<'PY'
import json
l=json.loads(open('gpurun_out/final_bench_4gpu.log').read().strip().splitlines()[-1])
print(l['parity_check']['estep'], l['parity_check']['mstep'], l['ms_per_iteration'], l['inversions_per_iteration'])
PY
