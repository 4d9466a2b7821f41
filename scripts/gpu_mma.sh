#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_row_split.py -m gpu -x -q > gpurun_out/mma_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/mma_pytest.log)"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), round(d["tok_s"]))'
for t in q4_K q4_0; do for n in 3 4 8; do
  for w in 8 16; do
  GGML_B200_MMA_WARPS=$w timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --type $t --ncols $n > gpurun_out/mma_${t}_${n}_${w}.log 2>&1
  echo "$t n=$n mma warps=$w: $(tail -1 gpurun_out/mma_${t}_${n}_${w}.log | python -c "$J" 2>&1 | tail -1)"
  done
done; done
GGML_B200_MMA_MIN=2 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --ncols 2 > gpurun_out/mma_n2.log 2>&1; echo "q4_K n=2 via mma: $(tail -1 gpurun_out/mma_n2.log | python -c "$J" 2>&1 | tail -1)"
GGML_B200_MMA_MIN=1 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --ncols 1 > gpurun_out/mma_n1.log 2>&1; echo "q4_K n=1 via mma: $(tail -1 gpurun_out/mma_n1.log | python -c "$J" 2>&1 | tail -1)"
