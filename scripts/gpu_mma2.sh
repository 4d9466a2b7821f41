#!/bin/bash
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), round(d["tok_s"]))'
for cfg in "0 0" "0 8" "0 12" "0 16"; do
  set -- $cfg
  GGML_B200_DBG=$1 GGML_B200_MMA_WARPS=$2 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --ncols 8 > gpurun_out/mma2_$1_$2.log 2>&1
  echo "q4_K n=8 dbg=$1 warps=$2: $(tail -1 gpurun_out/mma2_$1_$2.log | python -c "$J" 2>&1 | tail -1)"
done
GGML_B200_DBG=8 GGML_B200_MMA_WARPS=9 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mul_mat_vs_oracle or more_than_8 or full_size" 2>&1 | tail -1
