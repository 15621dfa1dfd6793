#!/bin/bash
# round 2, call k (2 GPUs): multi-GPU parity tests, 2-rank bench, the NCCL host program against the single-GPU run
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_engine_multigpu.py -m gpu -q > gpurun_out/r2k_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2k_tests.log
tail -5 gpurun_out/r2k_tests.log | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_bench2.log 2> gpurun_out/r2k_bench2.err; echo "bench2 rc $?"
tail -c 800 gpurun_out/r2k_bench2.err
python tools/show2.py gpurun_out/r2k_bench2.log
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r2k_bench2.log').read().strip().splitlines()[-1])
print(l['parity_check']); print(l['ms_per_iteration'], l['inversions_per_iteration'])
PY
make -C host cnf2freq_gpu_nccl > /dev/null
G=tests/golden/host
./host/cnf2freq_gpu --mapfile $G/f2.map --pedfile $G/f2.ped --genfile $G/f2.gen --count 4 --output gpurun_out/r2k_host1.out > gpurun_out/r2k_host1.stdout 2>/dev/null; echo "host1 rc $?"
rm -f /tmp/cnf2_nccl_id
for R in 0 1; do RANK=$R WORLD_SIZE=2 LOCAL_RANK=$R CNF2_NCCL_ID_FILE=/tmp/cnf2_nccl_id timeout 300 ./host/cnf2freq_gpu_nccl --mapfile $G/f2.map --pedfile $G/f2.ped --genfile $G/f2.gen --count 4 --output gpurun_out/r2k_host2_r$R.out > gpurun_out/r2k_host2_r$R.stdout 2> gpurun_out/r2k_host2_r$R.stderr & done; wait; echo "host2 done"
tail -n 3 gpurun_out/r2k_host2_r0.stderr; tail -n 3 gpurun_out/r2k_host2_r1.stderr
cmp gpurun_out/r2k_host1.out gpurun_out/r2k_host2_r0.out && echo "2-rank output identical to the 1-rank output"
diff <(grep -v Creating gpurun_out/r2k_host1.stdout) <(grep -v Creating gpurun_out/r2k_host2_r0.stdout) | head -10
