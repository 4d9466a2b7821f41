#!/bin/bash
# K7 (fused GLU / bias epilogues + GLU / ADD / ADD_ID nodes): parity tests, then the decode headline (the plain mat-vec must not move)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fusion_k7.py -m gpu -q 2>&1 | tail -15
timeout 900 python -m pytest tests/test_gpu_backend.py -m gpu -q -k "fused or k7" 2>&1 | tail -25
timeout 300 python bench.py --steps 20 --no-cpu > gpurun_out/k7_bench_default.json 2> gpurun_out/k7_bench_default.err
python -c "
import json; d=json.loads(open('gpurun_out/k7_bench_default.json').read().strip().splitlines()[-1]); print('decode q4_K n=1', round(d['value'],1), d['unit'], round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))"
timeout 300 python bench.py --steps 20 --no-cpu --fuse-glu > gpurun_out/k7_bench_fuse_glu.json 2> gpurun_out/k7_bench_fuse_glu.err
python -c "
import json; d=json.loads(open('gpurun_out/k7_bench_fuse_glu.json').read().strip().splitlines()[-1]); print('decode q4_K n=1 fused SWIGLU', round(d['value'],1), d['unit'], round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3), 'launches', d.get('gpu_launches'))"
