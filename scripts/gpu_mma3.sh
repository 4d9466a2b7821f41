#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_row_split.py -m gpu -x -q > gpurun_out/mma_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/mma_pytest.log)"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), round(d["tok_s"]))'
for t in q4_K q4_0; do for n in 3 4 5 8; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --type $t --ncols $n > gpurun_out/mma3_${t}_${n}.log 2>&1
  echo "$t n=$n: $(tail -1 gpurun_out/mma3_${t}_${n}.log | python -c "$J" 2>&1 | tail -1)"
done; done
