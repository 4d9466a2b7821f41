#!/bin/bash
# pass 8 (1 GPU): full GPU suite, smoke, default bench + reference arm, decode n = 1..8, all types, prefill, MoE shapes
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/p8_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p8_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/p8_smoke.log 2>&1; tail -1 gpurun_out/p8_smoke.log
timeout 600 python bench.py > gpurun_out/p8_bench_default.json 2> gpurun_out/p8_bench_default.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/p8_bench_reference.json 2> gpurun_out/p8_bench_reference.err; echo "reference arm rc=$?"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],3), "tok/s", round(d.get("tok_s",0)), "e2e", round(d.get("e2e",{}).get("value",0),1))'
echo "default: $(tail -1 gpurun_out/p8_bench_default.json | python -c "$J" 2>&1 | tail -1)"
for t in q4_K q4_0 q8_0; do for n in 1 2 3 4 8; do
  timeout 300 python bench.py --type $t --ncols $n --steps 10 --no-cpu > gpurun_out/p8_dec_${t}_n$n.json 2>/dev/null
  echo "decode $t n=$n: $(tail -1 gpurun_out/p8_dec_${t}_n$n.json | python -c "$J" 2>&1 | tail -1)"
done; done
timeout 300 python bench.py --workload 70b --steps 10 --no-cpu > gpurun_out/p8_dec_70b.json 2>/dev/null; echo "decode q4_K 70b: $(tail -1 gpurun_out/p8_dec_70b.json | python -c "$J" 2>&1 | tail -1)"
for t in q4_K q4_0 q8_0; do timeout 600 python bench.py --mode prefill --type $t --steps 5 --no-cpu > gpurun_out/p8_prefill_$t.json 2>/dev/null; echo "prefill $t: $(tail -1 gpurun_out/p8_prefill_$t.json | python -c "$J" 2>&1 | tail -1)"; done
timeout 300 python scripts/moe_bench.py > gpurun_out/p8_moe_q4_0.txt 2>&1; grep "tokens=" gpurun_out/p8_moe_q4_0.txt | head -14
