#!/bin/bash
# scripts/llama_ncu.sh TYPE LAYERS SKIP COUNT [ENV...] — per-kernel durations of one steady-state token step through libllama
# (oracle/_ref/llama_harness, whole graph on libggml-b200.so), after the same command has run clean without ncu.
set -u
T=${1:-q4_K}; L=${2:-32}; SKIP=${3:-4500}; CNT=${4:-700}; shift 4 || true
O=gpurun_out; mkdir -p $O
H=oracle/_ref/llama_harness
M=/tmp/ncu_model_$T
$H gen --type $T --out $M --layers $L --vocab 32000 --splits 1 > /dev/null || exit 1
ARGS="run --model $M --loader file --ngl 99 --prompt 16 --gen 12 --warmup 4 --fa --kqv --backend oracle/_ref/libggml-cpu-x64.so --backend ${LIB:-qvac-ext-lib-llama.cpp_b200/libggml-b200.so}"
env "$@" timeout 600 $H $ARGS > $O/llama_ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/llama_ncu_plain.log; exit 1; }
env "$@" timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip $SKIP -c $CNT --csv --log-file $O/llama_launches_${TAG:-$T}.csv $H $ARGS > $O/llama_ncu.log 2>&1
echo "ncu rc=$?"; wc -l $O/llama_launches_${TAG:-$T}.csv
