#!/bin/bash
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3))'
for n in 2 3 4; do
  for cfg in "0 0" "16 2" "12 2" "8 2"; do
    set -- $cfg
    GGML_B200_MV_WARPS=$1 GGML_B200_MV_STAGES=$2 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --ncols $n > gpurun_out/e3_${n}_${1}.log 2>&1
    echo "n=$n w=$1 s=$2: $(tail -1 gpurun_out/e3_${n}_${1}.log | python -c "$J" 2>&1 | tail -1)"
  done
done
timeout 200 python scripts/mv_timeline.py --layers 2 --ncols 4 --out gpurun_out/tl_n4.txt > /dev/null 2>&1
GGML_B200_MV_WARPS=8 GGML_B200_MV_STAGES=2 timeout 200 python scripts/mv_timeline.py --layers 2 --ncols 4 --out gpurun_out/tl_n4_w8.txt > /dev/null 2>&1
