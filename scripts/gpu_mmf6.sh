#!/bin/bash
# kind::f16 prefill GEMM after the two-phase fix-up + shared activation operand: correctness, timeline, schedule knobs
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python scripts/mmf_check.py > $O/mmf6_check.log 2>&1; echo "check rc=$? $(tail -1 $O/mmf6_check.log)"
# (pytest: see gpu_pass scripts)
timeout 240 python scripts/mmf_timeline.py --out $O/mmf6_timeline.txt > /dev/null 2>&1; cat $O/mmf6_timeline.txt
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), d["roofline"]["launches_per_step"])'
run() { echo "$1: $(env $2 timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type q4_K $3 2>/dev/null | tail -1 | python -c "$J" 2>&1 | tail -1)"; }
run "default" "X=1" ""

run "min_span=8" "GGML_B200_MMF_MIN_SPAN=8" ""
run "min_span=16" "GGML_B200_MMF_MIN_SPAN=16" ""

run "q8_0" "X=1" "--type q8_0"
