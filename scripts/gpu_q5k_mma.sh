#!/bin/bash
# q5_K on the tensor-core mat-vec kernel (3..8 columns, grouped MUL_MAT_ID)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "q5_K or 13" 2>&1 | tail -8
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],3), "tok/s", round(d.get("tok_s",0)))'
for n in 4 8; do
  timeout 300 python bench.py --type q5_K --ncols $n --steps 10 --no-cpu > gpurun_out/q5k_mma_n$n.json 2> gpurun_out/q5k_mma_n$n.err
  echo "decode q5_K n=$n: $(python -c "$J" < gpurun_out/q5k_mma_n$n.json 2>&1 | tail -1)"
done
