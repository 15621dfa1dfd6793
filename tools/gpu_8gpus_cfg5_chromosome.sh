#!/bin/bash
# One whole chromosome of BASELINE.json configs[4] (100,000 F2 in 500 families of 200 from 100 lines x 5,000 markers), dealt to 8 ranks
mkdir -p gpurun_out
nvidia-smi -L | wc -l; free -g | head -2
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8"
timeout 900 $R --workload cfg5chrom --strong --individuals 100000 --steps 3 --warmup 2 --no-cpu-baseline --no-extra-configs > gpurun_out/r2r_cfg5chrom100k_8.log 2> gpurun_out/r2r_cfg5chrom100k_8.err; echo "rc $?"
tail -n 4 gpurun_out/r2r_cfg5chrom100k_8.err | cut -c1-300
python tools/show2.py gpurun_out/r2r_cfg5chrom100k_8.log
