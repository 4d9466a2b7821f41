#!/bin/bash
# q6_K: decode n = 1, 2, 4 and prefill n = 512 on the Llama-3-8B stack (all matrices q6_K)
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],3), "tok/s", round(d.get("tok_s",0)))'
for n in 1 2 4; do
  timeout 300 python bench.py --type q6_K --ncols $n --steps 10 --no-cpu > gpurun_out/q6k_dec_n$n.json 2> gpurun_out/q6k_dec_n$n.err
  echo "decode q6_K n=$n: $(tail -1 gpurun_out/q6k_dec_n$n.json | python -c "$J" 2>&1 | tail -1)"
done
timeout 600 python bench.py --mode prefill --type q6_K --steps 5 --no-cpu > gpurun_out/q6k_prefill.json 2> gpurun_out/q6k_prefill.err
echo "prefill q6_K: $(tail -1 gpurun_out/q6k_prefill.json | python -c "$J" 2>&1 | tail -1)"
timeout 300 python bench.py --type q6_K --steps 5 --cpu-seconds 5 > gpurun_out/q6k_dec_cpu.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/q6k_dec_cpu.json').read().strip().splitlines()[-1]); print('cpu_baseline', d['cpu_baseline'])"
