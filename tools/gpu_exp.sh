cp cnf2freq_b200/lib/libcnf2_engine.so /tmp/orig.so
for v in 0 1 2 3 4; do
  cp tools/xlibs/libx$v.so cnf2freq_b200/lib/libcnf2_engine.so
  timeout 300 python bench.py --workload fullsib --individuals 1320 --chroms 2 --markers-per-chrom 200 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/exp_x$v.log 2>&1
done
cp /tmp/orig.so cnf2freq_b200/lib/libcnf2_engine.so
