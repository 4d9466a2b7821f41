#!/bin/bash
# scripts/gpu.sh — the one GPU-box runner (replaces the per-pass gpu_*.sh scripts of round 1). Runs on the box through
#   gpurun --timeout T -- 'bash scripts/gpu.sh <recipe> [<recipe> ...]'
# and leaves everything under gpurun_out/<tag>_*. Numbers printed under ncu are never bench values.
#   TAG=r02a bash scripts/gpu.sh tests greedy bench
# recipes:
#   tests          pytest -m gpu (everything except the llama greedy streams)
#   greedy         the greedy-decode token-stream parity test through libllama (3 types x 3 loaders)
#   bench          default bench.py line (+ reference arm line)
#   bench:ARGS     bench.py with ARGS (comma separated, e.g. bench:--ncols,4,--no-cpu); output tagged by a hash of ARGS
#   sweep:VAR:a,b,c:ARGS   bench.py once per value of environment variable VAR (decode experiments; light flags are added)
#   ncu-list:ARGS  ncu launch list (gpu__time_duration) of bench.py ARGS, after the same command has run clean without ncu
#   ncu-full:KERNEL_REGEX:ARGS  one ncu --set full capture (6 launches) of the kernels matching the regex
set -u
O=gpurun_out; mkdir -p $O
TAG=${TAG:-r02}
LIGHT="--no-cpu --no-ref-cuda --no-plugin-e2e --no-scale-base --no-token-step"
for recipe in "$@"; do
  kind=${recipe%%:*}; rest=${recipe#*:}; [ "$rest" == "$recipe" ] && rest=""
  case $kind in
    tests)
      timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_llama_greedy.py > $O/${TAG}_tests.log 2>&1
      echo "tests rc=$? $(tail -1 $O/${TAG}_tests.log)";;
    greedy)
      timeout 1500 python -m pytest tests/test_llama_greedy.py -m gpu -q > $O/${TAG}_greedy.log 2>&1
      echo "greedy rc=$? $(tail -1 $O/${TAG}_greedy.log)"; tail -30 $O/${TAG}_greedy.log | cut -c1-400;;
    bench)
      if [ -z "$rest" ]; then
        timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
        timeout 1500 python bench.py > $O/${TAG}_bench_default.json 2> $O/${TAG}_bench_default.err
        echo "bench rc=$?"; cut -c1-1500 $O/${TAG}_bench_default.json
      else
        args=${rest//,/ }; h=$(echo "$rest" | tr -c 'A-Za-z0-9_\n' '_' | cut -c1-60)
        timeout 1500 python bench.py $args > $O/${TAG}_bench_${h}.json 2> $O/${TAG}_bench_${h}.err
        echo "bench $args rc=$?"; cut -c1-600 $O/${TAG}_bench_${h}.json
      fi;;
    sweep)
      var=${rest%%:*}; r2=${rest#*:}; vals=${r2%%:*}; args=${r2#*:}; [ "$args" == "$r2" ] && args=""
      args=${args//,/ }
      for v in ${vals//,/ }; do
        env $var=$v timeout 600 python bench.py $LIGHT $args > $O/${TAG}_sweep_${var}_${v}.json 2> $O/${TAG}_sweep_${var}_${v}.err
        echo "$var=$v $args: $(python -c "import json,sys; j=json.load(open('$O/${TAG}_sweep_${var}_${v}.json')); print(round(j['value'],1), j['unit'], round(j['ms_per_step'],4), 'ms')" 2>&1 | tail -1)"
      done;;
    ncu-list)
      args=${rest//,/ }; h=$(echo "$rest" | tr -c 'A-Za-z0-9_\n' '_' | cut -c1-40)
      timeout 600 python bench.py $LIGHT $args > $O/${TAG}_plain_${h}.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches_${h}.csv python bench.py $LIGHT $args > $O/${TAG}_ncu_list_${h}.log 2>&1
      echo "ncu-list $args rc=$?";;
    ncu-full)
      rx=${rest%%:*}; args=${rest#*:}; args=${args//,/ }; h=$(echo "$rx" | tr -c 'A-Za-z0-9_\n' '_' | cut -c1-40)
      timeout 600 python bench.py $LIGHT $args > $O/${TAG}_plain_full_${h}.log 2>&1 &&
      timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$rx -s ${NCU_SKIP:-10} -c ${NCU_COUNT:-6} -o $O/${TAG}_${h} python bench.py $LIGHT $args > $O/${TAG}_ncu_full_${h}.log 2>&1
      echo "ncu-full $rx rc=$?";;
    *) echo "unknown recipe $recipe";;
  esac
done
