#!/bin/bash
# pass 9 (1 GPU): full GPU suite, smoke, default bench + reference arm, decode n = 1..8, prefill on both arithmetics,
# launch list + full capture of the kind::f16 prefill GEMM, full capture of the tensor-core mat-vec (n = 8)
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $O/p9_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/p9_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/p9_smoke.log 2>&1; tail -1 $O/p9_smoke.log
timeout 600 python bench.py > $O/p9_bench_default.json 2> $O/p9_bench_default.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/p9_bench_reference.json 2> $O/p9_bench_reference.err; echo "reference arm rc=$?"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],3), "tok/s", round(d.get("tok_s",0)), "e2e", round(d.get("e2e",{}).get("value",0),1))'
echo "default: $(tail -1 $O/p9_bench_default.json | python -c "$J" 2>&1 | tail -1)"
for n in 2 3 4 8; do
  timeout 300 python bench.py --type q4_K --ncols $n --steps 10 --no-cpu > $O/p9_dec_q4_K_n$n.json 2>/dev/null
  echo "decode q4_K n=$n: $(tail -1 $O/p9_dec_q4_K_n$n.json | python -c "$J" 2>&1 | tail -1)"
done
for t in q4_0 q8_0; do for n in 1 8; do
  timeout 300 python bench.py --type $t --ncols $n --steps 10 --no-cpu > $O/p9_dec_${t}_n$n.json 2>/dev/null
  echo "decode $t n=$n: $(tail -1 $O/p9_dec_${t}_n$n.json | python -c "$J" 2>&1 | tail -1)"
done; done
for t in q4_K q4_0 q8_0; do
  timeout 600 python bench.py --mode prefill --type $t --steps 5 --no-cpu > $O/p9_prefill_f16_$t.json 2>/dev/null
  echo "prefill f16 $t: $(tail -1 $O/p9_prefill_f16_$t.json | python -c "$J" 2>&1 | tail -1)"
done
timeout 600 python bench.py --mode prefill --arith i8 --type q4_K --steps 5 --no-cpu > $O/p9_prefill_i8_q4_K.json 2>/dev/null
echo "prefill i8 q4_K: $(tail -1 $O/p9_prefill_i8_q4_K.json | python -c "$J" 2>&1 | tail -1)"
timeout 240 python scripts/mmf_shapes.py --type q4_K --tokens 512 > $O/p9_mmf_shapes.log 2>&1; grep -v "^$" $O/p9_mmf_shapes.log | tail -12
# ---- profiler passes (numbers under ncu are never bench values)
P="python bench.py --mode prefill --type q4_K --steps 1 --warmup 3 --no-cpu --no-graph --layers 1"
timeout 300 $P > $O/p9_plain_prefill.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $O/p9_launches_prefill.csv $P > $O/p9_ncu_list.log 2>&1
echo "ncu prefill list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mmf_kernel -s 8 -c 4 -o $O/p9_mmf_prof $P > $O/p9_ncu_mmf.log 2>&1
echo "ncu mmf full rc=$?"
D="python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --layers 2 --ncols 8"
timeout 300 $D > $O/p9_plain_mma.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mv_mma -s 12 -c 4 -o $O/p9_mma_prof $D > $O/p9_ncu_mma.log 2>&1
echo "ncu mma full rc=$?"
