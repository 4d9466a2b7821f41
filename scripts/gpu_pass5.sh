#!/bin/bash
# pass 5 (1 GPU): cross-kernel weight prefetch
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "prefetch or golden or shared" > gpurun_out/p5_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p5_pytest.log
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), d.get("e2e",{}).get("value"))'
for kb in 0 32 64 128 256 512; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --prefetch-kb $kb $( [ $kb = 0 ] && echo --no-prefetch ) > gpurun_out/p5_pf$kb.log 2>&1
  echo "q4_K prefetch $kb KB: $(tail -1 gpurun_out/p5_pf$kb.log | python -c "$J" 2>&1)"
done
for t in q4_0 q8_0; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --type $t > gpurun_out/p5_$t.log 2>&1
  echo "$t prefetch 128: $(tail -1 gpurun_out/p5_$t.log | python -c "$J" 2>&1)"
done
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --ncols 2 > gpurun_out/p5_n2.log 2>&1; echo "n=2: $(tail -1 gpurun_out/p5_n2.log | python -c "$J" 2>&1)"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --workload 70b > gpurun_out/p5_70b.log 2>&1; echo "70b: $(tail -1 gpurun_out/p5_70b.log | python -c "$J" 2>&1)"
GGML_B200_MV_WARPS=8 GGML_B200_MV_STAGES=3 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/p5_w8.log 2>&1; echo "w8s3 pf128: $(tail -1 gpurun_out/p5_w8.log | python -c "$J" 2>&1)"
