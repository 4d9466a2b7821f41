#!/bin/bash
# row split over N GPUs, prefill (70B q4_K stack, n = 512): slice scatter kernel against peer stores from the GEMM epilogue
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac/GPU", round(d["roofline"]["frac"],3))'
N=${1:-2}
if [ "$N" = "2" ]; then timeout 300 python -m pytest tests/test_row_split.py -m gpu -x -q 2>&1 | tail -2; fi
PORT=29540
run() {
  PORT=$((PORT+1))
  env $3 timeout -k 10 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 5 --warmup 3 --no-cpu --mode prefill $2 > gpurun_out/sp_split${N}_$1.log 2>&1
  echo "split $N prefill $1 rc=$?: $(grep '^{' gpurun_out/sp_split${N}_$1.log | tail -1 | python -c "$J" 2>&1 | tail -1)"
  grep -i -E "note:|error|assert|Traceback|timed out" gpurun_out/sp_split${N}_$1.log | head -5
  grep '^{' gpurun_out/sp_split${N}_$1.log | tail -1 > gpurun_out/sp_split${N}_$1.json
}
run scatter "" "X=1"
if [ "${2:-}" = "ab" ]; then run epilogue "" "GGML_B200_MMF_SCATTER=0"; fi
