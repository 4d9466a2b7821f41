#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/p3_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p3_pytest.log
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), d.get("e2e",{}).get("value"))'
echo "--- decode q4_K variants"
for v in "X=1" "GGML_B200_MV_WARPS=16 GGML_B200_MV_STAGES=2" "GGML_B200_MV_WARPS=16 GGML_B200_MV_STAGES=3" "GGML_B200_MV_WARPS=12 GGML_B200_MV_STAGES=3" "GGML_B200_MV_WARPS=12 GGML_B200_MV_STAGES=4" "GGML_B200_MV_WARPS=8 GGML_B200_MV_STAGES=4"; do
  env $v timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/p3_var.log 2>&1
  echo "$v: $(tail -1 gpurun_out/p3_var.log | python -c "$J" 2>&1)"
done
for t in q4_0 q8_0; do for v in "X=1" "GGML_B200_MV_WARPS=16 GGML_B200_MV_STAGES=2" "GGML_B200_MV_WARPS=12 GGML_B200_MV_STAGES=2"; do
  env $v timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --type $t > gpurun_out/p3_var.log 2>&1
  echo "$t $v: $(tail -1 gpurun_out/p3_var.log | python -c "$J" 2>&1)"
done; done
echo "--- prefill"
for t in q8_0 q4_K; do timeout 600 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type $t > gpurun_out/p3_prefill_$t.log 2>&1; echo "prefill $t: $(tail -1 gpurun_out/p3_prefill_$t.log | python -c "$J" 2>&1)"; done
echo "--- ncu"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 > gpurun_out/p3_plain1.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 60 --csv --log-file gpurun_out/p3_launches_decode.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 > gpurun_out/p3_ncu1.log 2>&1
echo "ncu decode list rc=$?"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 > gpurun_out/p3_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mv_fast -s 24 -c 5 -o gpurun_out/p3_prof_mv python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 > gpurun_out/p3_ncu2.log 2>&1
echo "ncu mv full rc=$?"
timeout 300 python bench.py --mode prefill --type q8_0 --steps 1 --warmup 3 --no-cpu --no-graph --layers 1 > gpurun_out/p3_plain3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mmq_kernel -s 20 -c 2 -o gpurun_out/p3_prof_mmq python bench.py --mode prefill --type q8_0 --steps 1 --warmup 3 --no-cpu --no-graph --layers 1 > gpurun_out/p3_ncu3.log 2>&1
echo "ncu mmq full rc=$?"
