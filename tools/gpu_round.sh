set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t6_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/t6_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/t6_bench.log 2>&1; echo "bench rc $?" >> gpurun_out/t6_bench.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/t6_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/t6_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:warp_estep -c 1 -o gpurun_out/t6_warp_full python bench.py --individuals 1036 --chroms 1 --markers-per-chrom 1000 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/t6_ncu_full.log 2>&1
