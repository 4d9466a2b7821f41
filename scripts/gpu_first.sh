#!/bin/bash
# first GPU pass: parity tests, smoke, bench, launch list, one full ncu capture of the mat-vec kernel
set -u
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia-smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_q4K.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench_q4K.log
for t in q4_0 q8_0; do timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --type $t > gpurun_out/bench_$t.log 2>&1; tail -1 gpurun_out/bench_$t.log; done
for cfg in "4 2" "4 4" "8 2" "8 3" "8 4"; do set -- $cfg
  GGML_B200_MV_WARPS=$1 GGML_B200_MV_STAGES=$2 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/sweep_w$1_s$2.log 2>&1
  echo "w=$1 s=$2: $(tail -1 gpurun_out/sweep_w$1_s$2.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"])' 2>&1)"
done
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/plain_ncu.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 1000 -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/plain_ncu2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mv_fast -s 400 -c 8 -o gpurun_out/prof_mv python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
