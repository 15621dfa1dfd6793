#!/bin/bash
# cnf2_preprocess at the bench workload's size: ranked (default) and with the literal sums on every cell
mkdir -p gpurun_out
{
timeout 200 python tools/time_preprocess.py 2000 10 1000
timeout 300 python tools/time_preprocess.py 2000 10 1000 literal
} > gpurun_out/time_preprocess.txt 2>&1
cat gpurun_out/time_preprocess.txt | cut -c1-300
