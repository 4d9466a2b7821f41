#!/bin/bash
# tensor-core mat-vec with the activation quantization fused into the kernel (no quantizer launch)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_backend.py tests/test_row_split.py -m gpu -x -q > gpurun_out/mma5_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/mma5_pytest.log)"; grep -E "^(FAILED|ERROR)" gpurun_out/mma5_pytest.log | head
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],3), "tok/s", round(d.get("tok_s",0)), "launches", d["roofline"]["launches_per_step"])'
run() { echo "$1: $(env $2 timeout 300 python bench.py --steps 10 --no-cpu $3 2>/dev/null | tail -1 | python -c "$J" 2>&1 | tail -1)"; }
for n in 3 4 8; do
  run "n=$n fused<=64" "X=1" "--ncols $n"
  run "n=$n fused<=128" "GGML_B200_MMA_FUSE_UNITS=128" "--ncols $n"
  run "n=$n unfused" "GGML_B200_MMA_FUSE_UNITS=0" "--ncols $n"
done
run "n=2 IMMA fused" "GGML_B200_MMA_MIN=2" "--ncols 2"
run "n=4 q8_0 fused<=64" "X=1" "--ncols 4 --type q8_0"
run "n=4 q4_0 fused<=64" "X=1" "--ncols 4 --type q4_0"
