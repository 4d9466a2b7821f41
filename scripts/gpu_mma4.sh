#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_row_split.py tests/test_gpu_backend.py -m gpu -x -q > gpurun_out/mma_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/mma_pytest.log)"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), round(d["tok_s"]))'
for n in 3 4 8; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --type q8_0 --ncols $n > gpurun_out/mma4_q8_0_${n}.log 2>&1
  echo "q8_0 n=$n: $(tail -1 gpurun_out/mma4_q8_0_${n}.log | python -c "$J" 2>&1 | tail -1)"
done
timeout 300 python scripts/moe_bench.py --type q8_0 --batches 8,64 2>&1 | grep "tokens="
