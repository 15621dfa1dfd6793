#!/bin/bash
# round 2, 8 GPUs: the default bench line (weak scaling) and one cfg5-shaped chromosome strongly scaled
mkdir -p gpurun_out
nvidia-smi -L | wc -l
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8"
timeout 600 $R --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2q_weak8.log 2> gpurun_out/r2q_weak8.err; echo "weak rc $?"
timeout 600 $R --workload cfg5chrom --strong --steps 3 --warmup 2 --no-cpu-baseline --no-parity-check > gpurun_out/r2q_strong8.log 2> gpurun_out/r2q_strong8.err; echo "strong rc $?"
tail -c 600 gpurun_out/r2q_weak8.err | tail -n 5
python tools/show2.py gpurun_out/r2q_weak8.log gpurun_out/r2q_strong8.log
python - <<'PY'
import json
for f in ("weak8","strong8"):
    try:
        l=json.loads(open(f'gpurun_out/r2q_{f}.log').read().strip().splitlines()[-1])
        print(f, l['ms_per_iteration'], l['inversions_per_iteration'], l.get('parity_check',{}) and l['parity_check']['estep'], l['e2e']['ms_per_step'])
    except Exception as ex: print(f, ex)
PY
