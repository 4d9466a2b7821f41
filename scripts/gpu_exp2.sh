#!/bin/bash
# experiment: CTA shapes that let consecutive mat-vec kernels co-reside (warps x stages through the environment)
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3))'
for cfg in "16 2" "11 2" "10 2" "9 2" "8 2" "8 3"; do
  set -- $cfg
  export GGML_B200_MV_WARPS=$1 GGML_B200_MV_STAGES=$2
  timeout 200 python scripts/mv_timeline.py --layers 3 --out gpurun_out/tlc_$1_$2.txt > /dev/null 2>> gpurun_out/tl_err.txt
  for t in q4_K q4_0; do
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-prefetch --type $t > gpurun_out/expc_$1_$2_$t.log 2>&1
    echo "w=$1 s=$2 $t: $(tail -1 gpurun_out/expc_$1_$2_$t.log | python -c "$J" 2>&1)"
  done
done
tail -3 gpurun_out/tl_err.txt
