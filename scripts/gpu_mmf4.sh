#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 240 python scripts/mmf_check.py > gpurun_out/mmf4_check.log 2>&1
echo "check rc=$? $(tail -1 gpurun_out/mmf4_check.log)"
timeout 240 python scripts/mmf_shapes.py --type q4_K --tokens 512 2>&1 | tee gpurun_out/mmf4_shapes.log | grep -v "^$"
timeout 240 python scripts/mmf_timeline.py --out gpurun_out/mmf4_timeline.txt
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), d.get("e2e",{}).get("value"))'
for t in q4_K q8_0; do
  timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type $t > gpurun_out/mmf4_prefill_$t.log 2>&1
  echo "prefill $t: $(tail -1 gpurun_out/mmf4_prefill_$t.log | python -c "$J" 2>&1 | tail -1)"
done
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_row_split.py tests/test_gpu_backend.py -m gpu -x -q > gpurun_out/mmf4_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/mmf4_pytest.log)"
