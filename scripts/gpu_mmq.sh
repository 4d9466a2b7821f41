#!/bin/bash
# MMQ check: parity tests + prefill bench for the three types
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mmq or more_than_8" > gpurun_out/mmq_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/mmq_pytest.log)"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), d.get("e2e",{}).get("value"))'
for t in q8_0 q4_0 q4_K; do
  timeout 600 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type $t > gpurun_out/mmq_prefill_$t.log 2>&1
  echo "prefill $t: $(tail -1 gpurun_out/mmq_prefill_$t.log | python -c "$J" 2>&1)"
done
