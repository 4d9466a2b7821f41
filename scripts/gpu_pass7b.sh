#!/bin/bash
# pass 7b (1 GPU): ncu --set full of the decode kernel (its own call: one profiler per call)
set -u
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 > gpurun_out/p7_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mv_fast -s 24 -c 5 -o gpurun_out/p7_prof_mv python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 > gpurun_out/p7_ncu2.log 2>&1
echo "ncu mv full rc=$?"
