#!/bin/bash
# row split over N GPUs (70B q4_K stack): flagged-payload exchange against the flag protocol; decode n = 1 and n = 8; prefill
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac/GPU", round(d["roofline"]["frac"],3), d["config"]["workload"][-36:])'
N=${1:-2}
PORT=29520
run() {  # name, extra args
  PORT=$((PORT+1))
  timeout -k 10 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 10 --warmup 3 --no-cpu $2 > gpurun_out/ll_split${N}_$1.log 2>&1
  echo "split $N $1 rc=$?: $(grep '^{' gpurun_out/ll_split${N}_$1.log | tail -1 | python -c "$J" 2>&1 | tail -1)"
  grep -i -E "note:|error|assert|Traceback|timed out" gpurun_out/ll_split${N}_$1.log | head -5
  grep '^{' gpurun_out/ll_split${N}_$1.log | tail -1 > gpurun_out/ll_split${N}_$1.json
}
run ll ""
run flag "--no-ll"
run ll_n8 "--ncols 8"
run flag_n8 "--ncols 8 --no-ll"
if [ "${2:-}" = "prefill" ]; then run prefill "--mode prefill --steps 5"; fi
