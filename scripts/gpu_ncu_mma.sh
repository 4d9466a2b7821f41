#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 --ncols 8 > gpurun_out/nmma_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mv_mma -s 12 -c 5 -o gpurun_out/nmma_prof python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 --ncols 8 > gpurun_out/nmma_ncu.log 2>&1
echo "ncu rc=$?"
