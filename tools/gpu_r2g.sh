#!/bin/bash
mkdir -p gpurun_out
python tools/dbg_ns.py mstep_run_fullsib 2 4 34 39 2>&1 | tail -8
python tools/dbg_ns.py mstep_inv_demo 8 3 1 6 2>&1 | tail -8
timeout 900 python -m pytest tests/test_mstep_gpu.py tests/test_em_loop_gpu.py -m gpu -q > gpurun_out/r2g_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/r2g_tests.log
tail -8 gpurun_out/r2g_tests.log | cut -c1-200
timeout 600 python bench.py --no-cpu-baseline --no-parity-check --no-extra-configs --warmup 3 --steps 3 > gpurun_out/r2g_bench.log 2> gpurun_out/r2g_bench.err; echo "rc $?"
python tools/show2.py gpurun_out/r2g_bench.log
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r2g_bench.log').read().strip().splitlines()[-1])
print(l['ms_per_iteration'], l['roofline']['mstep']['gradient_evaluations_per_iteration'], l['roofline']['mstep']['solves_on_integral_free_steps_per_iteration'])
PY
