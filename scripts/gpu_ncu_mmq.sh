#!/bin/bash
set -u
mkdir -p gpurun_out
T=${1:-q4_K}; K=${2:-mmq_q4k_kernel}; S=${3:-8}; C=${4:-3}
timeout 300 python bench.py --mode prefill --type $T --steps 1 --warmup 3 --no-cpu --no-graph --layers 1 > gpurun_out/nm_plain_$T.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c $C -o gpurun_out/nm_prof_$T python bench.py --mode prefill --type $T --steps 1 --warmup 3 --no-cpu --no-graph --layers 1 > gpurun_out/nm_ncu_$T.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/nm_ncu_$T.log | cut -c1-300
