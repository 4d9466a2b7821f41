#!/bin/bash
# kind::f16 prefill GEMM: device timeline of one layer (graph replay) + schedule knobs on the whole 8B stack
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 240 python scripts/mmf_timeline.py --out $O/mmf5_timeline.txt > /dev/null 2>&1; cat $O/mmf5_timeline.txt
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4))'
run() { echo "$1: $(env $2 timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type q4_K 2>/dev/null | tail -1 | python -c "$J" 2>&1 | tail -1)"; }
run "default" "X=1"
run "streamk=0" "GGML_B200_MMF_STREAMK=0"
run "min_span=16" "GGML_B200_MMF_MIN_SPAN=16"
run "min_span=32" "GGML_B200_MMF_MIN_SPAN=32"
run "bn=128" "GGML_B200_MMF_BN=128"
run "pf=2" "GGML_B200_MMF_PF=2"
run "no pdl" "GGML_B200_PDL=0"
