#!/bin/bash
# pass 4b (2 GPUs): row split, fused all-gather against ncclAllGather
set -u
mkdir -p gpurun_out
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), d["config"]["workload"][-40:])'
N=${1:-2}
for g in fused nccl; do
  timeout -k 10 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --gather $g > gpurun_out/p4_split${N}_$g.log 2>&1
  echo "split $N $g rc=$?: $(grep '^{' gpurun_out/p4_split${N}_$g.log | tail -1 | python -c "$J" 2>&1)"
  grep -i -E "note:|error|assert|Traceback" gpurun_out/p4_split${N}_$g.log | head -5
done
timeout -k 10 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --gather fused --deferred > gpurun_out/p4_split${N}_fused_def.log 2>&1
echo "split $N fused deferred: $(grep "^{" gpurun_out/p4_split${N}_fused_def.log | tail -1 | python -c "$J" 2>&1)"
timeout -k 10 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu --gather fused --mode prefill > gpurun_out/p4_split${N}_prefill.log 2>&1
echo "split $N prefill fused rc=$?: $(grep '^{' gpurun_out/p4_split${N}_prefill.log | tail -1 | python -c "$J" 2>&1)"
grep -i -E "note:|error|assert|Traceback" gpurun_out/p4_split${N}_prefill.log | head -5
