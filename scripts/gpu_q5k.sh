#!/bin/bash
# q5_K: parity (C-ABI + plugin), then decode n = 1, 2, 4 and prefill n = 512 on the Llama-3-8B stack (all matrices q5_K)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_fusion_k7.py tests/test_gpu_backend.py -m gpu -q -k "q5_K or 13 or golden" 2>&1 | tail -12
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],3), "tok/s", round(d.get("tok_s",0)))'
for n in 1 2 4; do
  timeout 300 python bench.py --type q5_K --ncols $n --steps 10 --no-cpu > gpurun_out/q5k_dec_n$n.json 2> gpurun_out/q5k_dec_n$n.err
  echo "decode q5_K n=$n: $(python -c "$J" < gpurun_out/q5k_dec_n$n.json 2>&1 | tail -1)"
done
timeout 600 python bench.py --mode prefill --type q5_K --steps 5 --no-cpu > gpurun_out/q5k_prefill.json 2> gpurun_out/q5k_prefill.err
echo "prefill q5_K: $(python -c "$J" < gpurun_out/q5k_prefill.json 2>&1 | tail -1)"
