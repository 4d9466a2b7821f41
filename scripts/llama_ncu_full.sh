#!/bin/bash
# scripts/llama_ncu_full.sh — ncu --set full captures of the f3 kernels inside a real libllama run (4 layers of 8B geometry, 512-token
# prompt + 8 tokens), after the same command has run clean without ncu. Reports land in gpurun_out/llama_<kernel>.ncu-rep
set -u
O=gpurun_out; mkdir -p $O
H=oracle/_ref/llama_harness
M=/tmp/ncu_model_q4_K
$H gen --type q4_K --out $M --layers 4 --vocab 32000 --splits 1 > /dev/null || exit 1
ARGS="run --model $M --loader file --ngl 99 --prompt 512 --gen 8 --warmup 2 --fa --kqv --backend oracle/_ref/libggml-cpu-x64.so --backend qvac-ext-lib-llama.cpp_b200/libggml-b200.so"
timeout 600 $H $ARGS > $O/llama_ncu_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
for spec in "fa_tile_kernel:2:2" "fa_kernel:20:3" "mv_lean_kernel.*Lb1EEE:20:3" "attn_prep_kernel:20:2" "rms_norm_units_kernel:2:1"; do
  rx=${spec%%:*}; r=${spec#*:}; skip=${r%%:*}; cnt=${r#*:}
  name=$(echo "$rx" | tr -c 'A-Za-z0-9_\n' '_' | cut -c1-30)
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -f -o $O/llama_$name $H $ARGS > $O/llama_ncu_full_$name.log 2>&1
  echo "$rx rc=$?"
done
ls -la $O/*.ncu-rep
