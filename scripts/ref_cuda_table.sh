#!/bin/bash
# ours vs the reference's own ggml-cuda (built for sm_100 by oracle/Makefile) on the same MUL_MAT stacks, same box:
# decode n = 1, 2, 4, 8 and prefill n = 512 for q4_K / q4_0 / q8_0. Output: gpurun_out/${TAG}_refcuda_*.json
O=gpurun_out; mkdir -p $O; TAG=${TAG:-r02}
L="--no-cpu --no-ref-cuda --no-plugin-e2e --no-scale-base"
for t in q4_K q4_0 q8_0; do
  for n in 1 2 4 8; do
    [ "$t" != "q4_K" ] && [ "$n" != "1" ] && [ "$n" != "8" ] && continue
    python bench.py --type $t --ncols $n $L > $O/${TAG}_ours_dec_${t}_n$n.json 2>/dev/null
    python bench.py --impl reference_cuda --type $t --ncols $n --steps 10 > $O/${TAG}_refcuda_dec_${t}_n$n.json 2>/dev/null
  done
  python bench.py --mode prefill --type $t $L --steps 5 > $O/${TAG}_ours_prefill_f16_${t}.json 2>/dev/null
  python bench.py --mode prefill --arith i8 --type $t $L --steps 5 > $O/${TAG}_ours_prefill_i8_${t}.json 2>/dev/null
  python bench.py --impl reference_cuda --mode prefill --type $t --steps 5 > $O/${TAG}_refcuda_prefill_${t}.json 2>/dev/null
done
python - <<'PY'
import json, glob, os
tag = os.environ.get("TAG", "r02")
def val(f):
    try:
        j = json.load(open(f)); return round(j["value"], 1), j.get("unit", "")
    except Exception as e:
        return None, str(e)[:40]
for f in sorted(glob.glob(f"gpurun_out/{tag}_ours_*.json")):
    r = f.replace("_ours_", "_refcuda_").replace("_prefill_f16_", "_prefill_").replace("_prefill_i8_", "_prefill_")
    print(os.path.basename(f)[len(tag) + 6:-5], "ours", val(f), "reference-cuda", val(r))
PY
