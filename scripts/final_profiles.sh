#!/bin/bash
# the measurements profiles/README.md quotes, in one GPU-box run: TAG=r02 bash scripts/final_profiles.sh  ->  gpurun_out/${TAG}_final_*
O=gpurun_out; mkdir -p $O; TAG=${TAG:-r02}
L="--no-cpu --no-ref-cuda --no-plugin-e2e --no-scale-base"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_final_bench_reference.json 2> $O/${TAG}_final_bench_reference.err
timeout 1500 python bench.py > $O/${TAG}_final_bench_default.json 2> $O/${TAG}_final_bench_default.err
echo "bench rc=$?"
for n in 2 3 4 8; do timeout 300 python bench.py --ncols $n $L > $O/${TAG}_final_dec_q4_K_n$n.json 2>/dev/null; done
for t in q4_0 q8_0 q5_K q6_K; do timeout 300 python bench.py --type $t $L > $O/${TAG}_final_dec_${t}_n1.json 2>/dev/null; done
timeout 300 python bench.py --glu fused $L > $O/${TAG}_final_k7_glu_fused.json 2>/dev/null
python scripts/mv_timeline.py --layers 3 --prefetch-kb 64 --out $O/${TAG}_final_timeline.txt > /dev/null 2>&1
# DRAM traffic of the step's launches under CUDA-graph replay, caches NOT flushed between kernels (the roofline.traffic figure)
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none --graph-profiling node \
  -c 600 --csv --log-file $O/${TAG}_final_traffic_graph.csv python bench.py $L --steps 2 --warmup 3 > $O/${TAG}_final_traffic_graph.log 2>&1
echo "ncu traffic rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mv_lean -s 10 -c 6 -o $O/${TAG}_final_mv_lean python bench.py $L --layers 2 --no-graph --steps 2 --warmup 3 > $O/${TAG}_final_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $O/${TAG}_final_*
