#!/bin/bash
# E-step time per launch with one part of the per-marker work compiled out (results are wrong by construction; timing only).
# Libraries, built here before the call (they travel with the snapshot):
#   for d in NOTAB NOPAIRS NOREDUCE NOTRANSFER NOTURNS; do nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -shared \
#     -Xcompiler -fPIC -DCNF2_X_$d -o cnf2freq_b200/lib/abl/lib_$d.so cnf2freq_b200/csrc/cnf2_engine.cu; done
# usage: tools/gpu_ablations.sh NOTAB NOPAIRS ...
mkdir -p gpurun_out
for d in "$@"; do
  CNF2_ENGINE_LIB=$PWD/cnf2freq_b200/lib/abl/lib_$d.so timeout 300 python bench.py --no-cpu-baseline --no-extra-configs --no-parity-check --warmup 3 --steps 3 > gpurun_out/abl_$d.json 2> gpurun_out/abl_$d.err
  python - "$d" <<'PY'
import json, sys
d = sys.argv[1]
try:
    j = json.loads(open(f"gpurun_out/abl_{d}.json").read().strip().splitlines()[-1])
    r = j["roofline"]
    print(f"{d}: estep {r['launch_ms']:.2f} ms/launch, isolated {r['launch_ms_isolated']:.2f}, iteration {j['ms_per_step']:.1f} ms")
except Exception as e:
    print(d, "failed", e)
PY
done
