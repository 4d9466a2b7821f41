#!/bin/bash
# first pass of the kind::f16 prefill GEMM: correctness per tile shape / cluster size, then the prefill bench
set -u
mkdir -p gpurun_out
for cfg in "128 1" "256 1" "128 2" "256 2" "0 0"; do
  set -- $cfg
  GGML_B200_MMF_BN=$1 GGML_B200_MMF_CL=$2 timeout 240 python scripts/mmf_check.py > gpurun_out/mmf_check_$1_$2.log 2>&1
  echo "check BN=$1 CL=$2 rc=$? $(tail -1 gpurun_out/mmf_check_$1_$2.log)"
done
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), d.get("e2e",{}).get("value"))'
for cfg in "256 1" "256 2" "128 2" "0 0"; do
  set -- $cfg
  for t in q4_K q8_0; do
    GGML_B200_MMF_BN=$1 GGML_B200_MMF_CL=$2 timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type $t > gpurun_out/mmf_prefill_${t}_$1_$2.log 2>&1
    echo "prefill $t BN=$1 CL=$2: $(tail -1 gpurun_out/mmf_prefill_${t}_$1_$2.log | python -c "$J" 2>&1 | tail -1)"
  done
done
GGML_B200_MMQ_F16=0 timeout 300 python bench.py --mode prefill --steps 5 --warmup 3 --no-cpu --type q4_K > gpurun_out/mmf_prefill_i8.log 2>&1
echo "prefill q4_K i8: $(tail -1 gpurun_out/mmf_prefill_i8.log | python -c "$J" 2>&1 | tail -1)"
# the PDL quantizer in front of the tensor-core mat-vec
J2='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), round(d["tok_s"]))'
for t in q4_K q8_0; do for n in 4 8; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --type $t --ncols $n > gpurun_out/mmf_dec_${t}_$n.log 2>&1
  echo "$t n=$n: $(tail -1 gpurun_out/mmf_dec_${t}_$n.log | python -c "$J2" 2>&1 | tail -1)"
done; done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "not mmq and not more_than_8" > gpurun_out/mmf_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/mmf_pytest.log)"
