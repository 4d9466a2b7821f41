#!/bin/bash
set -u
mkdir -p gpurun_out
for cfg in "256 1" "256 2" "128 1" "128 2"; do
  set -- $cfg
  echo "== BN=$1 CL=$2"
  GGML_B200_MMF_BN=$1 GGML_B200_MMF_CL=$2 timeout 240 python scripts/mmf_shapes.py --type q4_K --tokens 512,2048 2>&1 | tee gpurun_out/mmf_shapes_$1_$2.log | grep -v "^$"
done
echo "== i8"
timeout 240 python scripts/mmf_shapes.py --type q4_K --tokens 512 --i8 2>&1 | tee gpurun_out/mmf_shapes_i8.log
timeout 300 python bench.py --mode prefill --steps 2 --warmup 1 --no-cpu --type q4_K --layers 1 --no-graph > gpurun_out/mmf_l1.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/mmf_launches.csv python bench.py --mode prefill --steps 2 --warmup 1 --no-cpu --type q4_K --layers 1 --no-graph > gpurun_out/mmf_ncu.log 2>&1
echo "ncu rc=$?"
