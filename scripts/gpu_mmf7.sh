#!/bin/bash
# kind::f16 prefill GEMM with grouped (multi-matrix) launches: correctness, then the 8B stack with and without grouping
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python scripts/mmf_check.py > $O/mmf7_check.log 2>&1; echo "check rc=$? $(tail -1 $O/mmf7_check.log)"
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "shared or mmq or prefill or full_size" > $O/mmf7_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/mmf7_pytest.log)"; grep -E "^(FAILED|ERROR)|Error" $O/mmf7_pytest.log | head -5
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), d["roofline"]["launches_per_step"])'
run() { echo "$1: $(env $2 timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type q4_K $3 2>&1 | tail -1 | python -c "$J" 2>&1 | tail -1)"; }
run "grouped" "X=1" ""
run "group=0" "GGML_B200_MMF_GROUP=0" ""
run "grouped min_span=4" "GGML_B200_MMF_MIN_SPAN=4" ""
run "grouped min_span=16" "GGML_B200_MMF_MIN_SPAN=16" ""
run "grouped q8_0" "X=1" "--type q8_0"
