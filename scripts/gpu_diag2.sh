#!/bin/bash
# 2-GPU diagnostic: where does the time of a row-split step go (no exchange / fused flags / ncclAllGather)?
set -u
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/d2_topo.txt 2>&1; nvidia-smi nvlink --status 2>&1 | head -30 >> gpurun_out/d2_topo.txt
head -12 gpurun_out/d2_topo.txt
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), d["gpu_launches"]//d["steps"])'
N=2
run() { # name, extra args...
  name=$1; shift
  timeout -k 10 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu --layers 8 "$@" > gpurun_out/d2_$name.log 2>&1
  echo "$name rc=$?: $(grep '^{' gpurun_out/d2_$name.log | tail -1 | python -c "$J" 2>&1)"
  grep -i -E "note:|error|assert" gpurun_out/d2_$name.log | head -3
}
run none --gather none
run none_nopf --gather none --no-prefetch
run fused --gather fused
run fused_nopdl --gather fused
NCCL_DEBUG=INFO run nccl --gather nccl
grep -E "via|NVLS|P2P|SHM|Channel 00" gpurun_out/d2_nccl.log | head -8
GGML_B200_PDL=0 run fused_pdl0 --gather fused
run fused_nograph --gather fused --no-graph
timeout 120 python bench.py --steps 5 --warmup 3 --no-cpu --layers 8 --workload 70b > gpurun_out/d2_single.log 2>&1; echo "single: $(tail -1 gpurun_out/d2_single.log | python -c "$J" 2>&1)"
