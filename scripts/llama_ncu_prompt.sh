#!/bin/bash
# scripts/llama_ncu_prompt.sh TYPE LAYERS PROMPT COUNT — per-kernel durations of a PROMPT pass through libllama (first COUNT launches
# after one untimed pass), whole graph on libggml-b200.so
set -u
T=${1:-q4_K}; L=${2:-4}; P=${3:-512}; CNT=${4:-400}
O=gpurun_out; mkdir -p $O
H=oracle/_ref/llama_harness
M=/tmp/ncu_model_$T
$H gen --type $T --out $M --layers $L --vocab 32000 --splits 1 > /dev/null || exit 1
ARGS="run --model $M --loader file --ngl 99 --prompt $P --gen 2 --fa --kqv --backend oracle/_ref/libggml-cpu-x64.so --backend ${LIB:-qvac-ext-lib-llama.cpp_b200/libggml-b200.so}"
timeout 600 $H $ARGS > $O/llama_ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/llama_ncu_plain.log; exit 1; }
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c $CNT --csv --log-file $O/llama_launches_prompt_${TAG:-$T}.csv $H $ARGS > $O/llama_ncu.log 2>&1
echo "ncu rc=$?"; wc -l $O/llama_launches_prompt_${TAG:-$T}.csv
