#!/bin/bash
# kind::f16 prefill GEMM on CTA pairs (cta_group::2, GGML_B200_MMF_CL=2): correctness first (short timeouts: a wrong barrier hangs), then speed
set -u
mkdir -p gpurun_out
O=gpurun_out
GGML_B200_MMF_CL=2 timeout 120 python scripts/mmf_check.py > $O/mmf8_check.log 2>&1; rc=$?; echo "check CL=2 rc=$rc $(tail -1 $O/mmf8_check.log)"
if [ $rc -ne 0 ]; then tail -5 $O/mmf8_check.log; nvidia-smi --query-gpu=name,memory.used --format=csv,noheader; exit 0; fi
grep -c nmse $O/mmf8_check.log; sort -t' ' -k2 $O/mmf8_check.log | grep nmse | awk '{print $NF}' | sort -g | tail -3
GGML_B200_MMF_CL=2 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "shared or mmq or prefill or full_size" > $O/mmf8_pytest.log 2>&1; echo "pytest CL=2 rc=$? $(tail -1 $O/mmf8_pytest.log)"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), d["roofline"]["launches_per_step"])'
run() { echo "$1: $(env $2 timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type q4_K $3 2>&1 | tail -1 | python -c "$J" 2>&1 | tail -1)"; }
run "CL=1" "X=1" ""
run "CL=2" "GGML_B200_MMF_CL=2" ""
run "CL=2 min_span=16" "GGML_B200_MMF_CL=2 GGML_B200_MMF_MIN_SPAN=16" ""
GGML_B200_MMF_CL=2 timeout 240 python scripts/mmf_shapes.py --type q4_K --tokens 512 > $O/mmf8_shapes.log 2>&1; grep -v "^$" $O/mmf8_shapes.log | tail -7
