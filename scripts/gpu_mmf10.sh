#!/bin/bash
# kind::f16 prefill GEMM, fix-up parts staged through shared memory by bulk copies: correctness (CL = 2 and 1), speed, timeline
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 120 python scripts/mmf_check.py > $O/mmf10_check.log 2>&1; rc=$?; echo "check CL=2 rc=$rc $(tail -1 $O/mmf10_check.log)"
if [ $rc -ne 0 ]; then tail -5 $O/mmf10_check.log; exit 0; fi
GGML_B200_MMF_CL=1 timeout 120 python scripts/mmf_check.py > $O/mmf10_check1.log 2>&1; echo "check CL=1 rc=$? $(tail -1 $O/mmf10_check1.log)"
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_row_split.py -m gpu -x -q -k "shared or mmq or prefill or full_size or exchange" > $O/mmf10_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/mmf10_pytest.log)"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), d["roofline"]["launches_per_step"])'
run() { echo "$1: $(env $2 timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type q4_K $3 2>&1 | tail -1 | python -c "$J" 2>&1 | tail -1)"; }
run "CL=2" "X=1" ""
run "CL=2 min_span=4" "GGML_B200_MMF_MIN_SPAN=4" ""
run "CL=1" "GGML_B200_MMF_CL=1" ""
timeout 240 python scripts/mmf_timeline.py --out $O/mmf10_timeline.txt > /dev/null 2>&1; cat $O/mmf10_timeline.txt
