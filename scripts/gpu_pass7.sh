#!/bin/bash
# pass 7 (1 GPU): full GPU suite, default bench line, reference arm, launch list + full capture of the decode kernel
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p7_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p7_pytest.log
timeout 600 python bench.py > gpurun_out/p7_bench_default.json 2> gpurun_out/p7_bench_default.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/p7_bench_default.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/p7_bench_reference.json 2> gpurun_out/p7_bench_reference.err; echo "reference arm rc=$?"; cut -c1-400 gpurun_out/p7_bench_reference.json
for t in q4_0 q8_0; do timeout 300 python bench.py --type $t --no-cpu > gpurun_out/p7_bench_$t.json 2>/dev/null; done
for t in q4_K q4_0 q8_0; do timeout 600 python bench.py --mode prefill --type $t --steps 5 --no-cpu > gpurun_out/p7_prefill_$t.json 2>/dev/null; done
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/p7_smoke.log 2>&1; tail -1 gpurun_out/p7_smoke.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 > gpurun_out/p7_plain1.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 60 --csv --log-file gpurun_out/p7_launches_decode.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 > gpurun_out/p7_ncu1.log 2>&1
echo "ncu decode list rc=$?"
