#!/bin/bash
# pass 12 profiler evidence (numbers under ncu are never bench values): q6_K decode kernel and the K7 fused gate/up/SWIGLU launch
set -u
O=gpurun_out; mkdir -p $O
Q="python bench.py --type q6_K --steps 2 --warmup 3 --no-cpu --no-graph --layers 2"
timeout 300 $Q > $O/p12_plain_q6k.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $O/p12_launches_q6k.csv $Q > $O/p12_ncu_list_q6k.log 2>&1
echo "ncu q6_K list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mv_fast -s 10 -c 6 -o $O/p12_q6k_mv $Q > $O/p12_ncu_q6k.log 2>&1
echo "ncu q6_K full rc=$?"
G="python bench.py --glu fused --steps 2 --warmup 3 --no-cpu --no-graph --layers 2"
timeout 300 $G > $O/p12_plain_glu.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mv_fast -s 10 -c 6 -o $O/p12_glu_mv $G > $O/p12_ncu_glu.log 2>&1
echo "ncu glu full rc=$?"
ls -la $O/*.ncu-rep
