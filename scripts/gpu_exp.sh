#!/bin/bash
# experiment runner: timeline + bench under a GGML_B200_DBG value
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3))'
for d in "$@"; do
  export GGML_B200_DBG=$d
  timeout 200 python scripts/mv_timeline.py --layers 3 --out gpurun_out/tlx_$d.txt > /dev/null 2>> gpurun_out/tl_err.txt
  timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mul_mat and not mmq" > gpurun_out/exp_pytest_$d.log 2>&1; echo "dbg=$d pytest rc=$? $(tail -1 gpurun_out/exp_pytest_$d.log)"
  for t in q4_K q8_0; do
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-prefetch --type $t > gpurun_out/exp_${d}_$t.log 2>&1
    echo "dbg=$d $t: $(tail -1 gpurun_out/exp_${d}_$t.log | python -c "$J" 2>&1)"
  done
done
tail -3 gpurun_out/tl_err.txt
