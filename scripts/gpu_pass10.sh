#!/bin/bash
# pass 10 (1 GPU): full GPU suite, smoke, default bench + reference arm, decode n = 1..8, prefill on both arithmetics,
# launch list + full capture of the grouped 2-CTA kind::f16 prefill GEMM
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $O/p10_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/p10_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/p10_smoke.log 2>&1; tail -1 $O/p10_smoke.log
timeout 600 python bench.py > $O/p10_bench_default.json 2> $O/p10_bench_default.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/p10_bench_reference.json 2> $O/p10_bench_reference.err; echo "reference arm rc=$?"
J='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), d["unit"], round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],3), "tok/s", round(d.get("tok_s",0)), "e2e", round(d.get("e2e",{}).get("value",0),1))'
echo "default: $(tail -1 $O/p10_bench_default.json | python -c "$J" 2>&1 | tail -1)"
for n in 2 4 8; do
  timeout 300 python bench.py --type q4_K --ncols $n --steps 10 --no-cpu > $O/p10_dec_q4_K_n$n.json 2>/dev/null
  echo "decode q4_K n=$n: $(tail -1 $O/p10_dec_q4_K_n$n.json | python -c "$J" 2>&1 | tail -1)"
done
echo "decode n=1 unfused quant, dp4a kernel: $(GGML_B200_FUSE_QUANT=0 timeout 300 python bench.py --steps 10 --no-cpu 2>/dev/null | tail -1 | python -c "$J" 2>&1 | tail -1)"
echo "decode n=1 unfused quant, IMMA kernel: $(GGML_B200_FUSE_QUANT=0 GGML_B200_MMA_MIN=1 timeout 300 python bench.py --steps 10 --no-cpu 2>/dev/null | tail -1 | python -c "$J" 2>&1 | tail -1)"
for t in q4_K q4_0 q8_0; do
  timeout 600 python bench.py --mode prefill --type $t --steps 5 --no-cpu > $O/p10_prefill_f16_$t.json 2>/dev/null
  echo "prefill f16 $t: $(tail -1 $O/p10_prefill_f16_$t.json | python -c "$J" 2>&1 | tail -1)"
done
timeout 600 python bench.py --mode prefill --arith i8 --type q4_K --steps 5 --no-cpu > $O/p10_prefill_i8_q4_K.json 2>/dev/null
echo "prefill i8 q4_K: $(tail -1 $O/p10_prefill_i8_q4_K.json | python -c "$J" 2>&1 | tail -1)"
# ---- profiler passes (numbers under ncu are never bench values)
P="python bench.py --mode prefill --type q4_K --steps 1 --warmup 3 --no-cpu --no-graph --layers 1"
timeout 300 $P > $O/p10_plain_prefill.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file $O/p10_launches_prefill.csv $P > $O/p10_ncu_list.log 2>&1
echo "ncu prefill list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mmf_kernel -s 5 -c 5 -o $O/p10_mmf_prof $P > $O/p10_ncu_mmf.log 2>&1
echo "ncu mmf full rc=$?"
