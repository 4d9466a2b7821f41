#!/bin/bash
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), d["roofline"]["launches_per_step"])'
run() { echo "$1: $(env $2 timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type q4_K $3 2>&1 | tail -1 | python -c "$J" 2>&1 | tail -1)"; }
run "act reg<=64" "X=1" ""
run "act reg<=32" "GGML_B200_MMF_ACT_REG=32" ""
run "act reg<=16" "GGML_B200_MMF_ACT_REG=16" ""
run "act reg<=0" "GGML_B200_MMF_ACT_REG=0" ""
