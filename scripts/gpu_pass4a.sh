#!/bin/bash
# pass 4a (1 GPU): full GPU test suite, decode bench with the current defaults, small-batch decode
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/p4_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p4_pytest.log
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), d.get("e2e",{}).get("value"))'
for t in q4_K q4_0 q8_0; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --type $t > gpurun_out/p4_dec_$t.log 2>&1
  echo "decode $t: $(tail -1 gpurun_out/p4_dec_$t.log | python -c "$J" 2>&1)"
done
for n in 2 4 8; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --ncols $n > gpurun_out/p4_dec_n$n.log 2>&1
  echo "decode q4_K n=$n: $(tail -1 gpurun_out/p4_dec_n$n.log | python -c "$J" 2>&1)"
done
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --workload 70b > gpurun_out/p4_dec_70b.log 2>&1
echo "decode q4_K 70b 1gpu: $(tail -1 gpurun_out/p4_dec_70b.log | python -c "$J" 2>&1)"
timeout 600 python bench.py > gpurun_out/p4_bench_default.log 2>&1; echo "default bench rc=$?"; tail -1 gpurun_out/p4_bench_default.log | cut -c1-1500
