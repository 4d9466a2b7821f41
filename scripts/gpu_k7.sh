#!/bin/bash
# K7 (fused GLU / bias epilogues + GLU / ADD / ADD_ID nodes): parity tests, then the decode headline (the plain mat-vec must not move)
# and the step with the SWIGLU node behind gate/up: element-wise launch vs the fused epilogue
set -u
mkdir -p gpurun_out
if [ "${1:-}" != "bench" ]; then
  timeout 900 python -m pytest tests/test_fusion_k7.py -m gpu -q 2>&1 | tail -15
  timeout 900 python -m pytest tests/test_gpu_backend.py -m gpu -q -k "fused or k7" 2>&1 | tail -25
fi
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],3), "launches", d.get("gpu_launches"))'
for g in none separate fused; do
  timeout 300 python bench.py --steps 20 --no-cpu --glu $g > gpurun_out/k7_bench_glu_$g.json 2> gpurun_out/k7_bench_glu_$g.err
  echo "decode q4_K n=1 glu=$g: $(python -c "$J" < gpurun_out/k7_bench_glu_$g.json 2>&1 | tail -1)"
done
