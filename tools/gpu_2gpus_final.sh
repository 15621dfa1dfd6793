#!/bin/bash
# 2 GPUs at the round's last state: the multi-GPU parity tests, a 2-rank bench line, the NCCL host program against the 1-rank run
mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 300 python -m pytest tests/test_engine_multigpu.py -m gpu -q --timeout 150 > gpurun_out/final_tests_2gpu.log 2>&1; echo "tests rc $?" >> gpurun_out/final_tests_2gpu.log
tail -4 gpurun_out/final_tests_2gpu.log | cut -c1-300
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_2gpu.log 2> gpurun_out/final_bench_2gpu.err; echo "bench2 rc $?"
tail -c 600 gpurun_out/final_bench_2gpu.err
python tools/show2.py gpurun_out/final_bench_2gpu.log
python - <<'PY'
import json
l=json.loads(open('gpurun_out/final_bench_2gpu.log').read().strip().splitlines()[-1])
print(l['parity_check']['estep'], l['parity_check']['mstep'], l['ms_per_iteration'], l['inversions_per_iteration'])
PY
G=tests/golden/host
./host/cnf2freq_gpu --mapfile $G/f2.map --pedfile $G/f2.ped --genfile $G/f2.gen --count 4 --output gpurun_out/host1.out > gpurun_out/host1.stdout 2>/dev/null; echo "host1 rc $?"
rm -f /tmp/cnf2_nccl_id
for R in 0 1; do RANK=$R WORLD_SIZE=2 LOCAL_RANK=$R CNF2_NCCL_ID_FILE=/tmp/cnf2_nccl_id timeout 200 ./host/cnf2freq_gpu_nccl --mapfile $G/f2.map --pedfile $G/f2.ped --genfile $G/f2.gen --count 4 --output gpurun_out/host2_r$R.out > gpurun_out/host2_r$R.stdout 2> gpurun_out/host2_r$R.stderr & done; wait; echo "host2 done"
tail -n 2 gpurun_out/host2_r0.stderr; tail -n 2 gpurun_out/host2_r1.stderr
cmp gpurun_out/host1.out gpurun_out/host2_r0.out && echo "2-rank output identical to the 1-rank output"
