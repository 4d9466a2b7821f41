#!/bin/bash
set -u
mkdir -p gpurun_out
for cfg in "1 0" "0 0" "1 8"; do
  set -- $cfg
  GGML_B200_MMF_STREAMK=$1 GGML_B200_MMF_PF=$2 timeout 240 python scripts/mmf_check.py > gpurun_out/mmf3_check_$1_$2.log 2>&1
  echo "check SK=$1 PF=$2 rc=$? $(tail -1 gpurun_out/mmf3_check_$1_$2.log)"
done
for cfg in "1 0" "0 0" "1 8" "1 16"; do
  set -- $cfg
  echo "== SK=$1 PF=$2"
  GGML_B200_MMF_STREAMK=$1 GGML_B200_MMF_PF=$2 timeout 240 python scripts/mmf_shapes.py --type q4_K --tokens 512 2>&1 | tee gpurun_out/mmf3_shapes_$1_$2.log | grep -v "^$"
done
GGML_B200_MMF_STREAMK=1 timeout 240 python scripts/mmf_timeline.py --out gpurun_out/mmf3_timeline_sk.txt
GGML_B200_MMF_STREAMK=0 timeout 240 python scripts/mmf_timeline.py --out gpurun_out/mmf3_timeline_nosk.txt
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), d.get("e2e",{}).get("value"))'
for t in q4_K q4_0 q8_0; do
  timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type $t > gpurun_out/mmf3_prefill_$t.log 2>&1
  echo "prefill $t: $(tail -1 gpurun_out/mmf3_prefill_$t.log | python -c "$J" 2>&1 | tail -1)"
done
