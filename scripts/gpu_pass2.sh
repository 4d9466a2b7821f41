#!/bin/bash
# second GPU pass: new warp-specialised mat-vec kernel (PDL, fused quantization) + first tcgen05 MMQ kernel
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "not mmq" > gpurun_out/p2_pytest_mv.log 2>&1; echo "pytest mv rc=$?"; tail -3 gpurun_out/p2_pytest_mv.log
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "mmq" > gpurun_out/p2_pytest_mmq.log 2>&1; echo "pytest mmq rc=$?"; tail -15 gpurun_out/p2_pytest_mmq.log
timeout 600 python -m pytest tests/test_gpu_backend.py -m gpu -x -q > gpurun_out/p2_pytest_backend.log 2>&1; echo "pytest backend rc=$?"; tail -3 gpurun_out/p2_pytest_backend.log
for t in q4_K q4_0 q8_0; do timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --type $t > gpurun_out/p2_bench_$t.log 2>&1; echo "bench $t rc=$?"; tail -1 gpurun_out/p2_bench_$t.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["roofline"]["frac"], d.get("e2e",{}).get("value"))'; done
echo "--- variants (q4_K)"
for v in "GGML_B200_PDL=0" "GGML_B200_FUSE_QUANT=0" "GGML_B200_MV_WARPS=8 GGML_B200_MV_STAGES=2" "GGML_B200_MV_WARPS=8 GGML_B200_MV_STAGES=4" "GGML_B200_MV_WARPS=6 GGML_B200_MV_STAGES=3" "GGML_B200_MV_WARPS=4 GGML_B200_MV_STAGES=4"; do
  env $v timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/p2_var.log 2>&1
  echo "$v: $(tail -1 gpurun_out/p2_var.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"])' 2>&1)"
done
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-graph > gpurun_out/p2_nograph.log 2>&1; echo "no-graph: $(tail -1 gpurun_out/p2_nograph.log | cut -c1-200)"
for n in 2 4 8; do timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --ncols $n > gpurun_out/p2_n$n.log 2>&1; echo "ncols=$n: $(tail -1 gpurun_out/p2_n$n.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["tok_s"])' 2>&1)"; done
echo "--- prefill"
for t in q8_0 q4_0 q4_K; do timeout 600 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type $t > gpurun_out/p2_prefill_$t.log 2>&1; echo "prefill $t rc=$?"; tail -1 gpurun_out/p2_prefill_$t.log | cut -c1-300; done
