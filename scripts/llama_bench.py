#!/usr/bin/env python
"""Decode / prefill speed through the REAL libllama (oracle/_ref/llama_harness, the reference's own graph builder, scheduler and KV
cache) on a random-init Llama-3-8B-geometry GGUF: libggml-b200.so with the whole graph on the device (f3 ops, offload_kqv, flash
attention) against the reference's CUDA backend built for sm_100 (oracle/Makefile target `cuda`) on the same file and flags.
What llama-bench reports as pp / tg, with its warm-up run. Runs on the GPU box:

    python scripts/llama_bench.py --type q4_K --layers 32 --prompt 512 --gen 128 --out gpurun_out/llama_bench_q4_K.json
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
REF = ROOT / "oracle" / "_ref"
HARNESS = REF / "llama_harness"
B200 = ROOT / "qvac-ext-lib-llama.cpp_b200" / "libggml-b200.so"


def run(args, env=None, timeout=1800):
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run([str(HARNESS)] + [str(a) for a in args], capture_output=True, text=True, timeout=timeout, env=e)
    if p.returncode != 0:
        return {"error": (p.stderr or p.stdout)[-1500:]}
    r = json.loads(p.stdout.strip().splitlines()[-1])
    t = [ln for ln in p.stderr.splitlines() if "ggml-b200 timing" in ln]   # GGML_B200_TIMING=1
    if t:
        r["backend_timing"] = t
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="q4_K")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--vocab", type=int, default=32000)
    ap.add_argument("--prompt", type=int, default=512)
    ap.add_argument("--gen", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--env", action="append", default=[], help="extra b200 configuration: NAME:K=V[,K=V...]")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    tmp = Path(tempfile.mkdtemp(prefix="llama_bench_"))
    prefix = tmp / f"m_{a.type}"
    g = run(["gen", "--type", a.type, "--out", prefix, "--layers", a.layers, "--vocab", a.vocab, "--splits", 1])
    if "error" in g:
        print(g["error"], file=sys.stderr)
        return 1
    common = ["run", "--model", prefix, "--loader", "file", "--ngl", 99, "--prompt", a.prompt, "--gen", a.gen, "--warmup", a.warmup, "--fa", "--kqv",
              "--backend", REF / "libggml-cpu-x64.so"]
    configs = [("b200", B200, {})]
    for spec in a.env:
        name, kv = spec.split(":", 1)
        configs.append((name, B200, dict(x.split("=", 1) for x in kv.split(","))))
    if not a.no_ref_cuda and (REF / "libggml-cuda.so").exists():
        configs.append(("reference_cuda", REF / "libggml-cuda.so", {}))
    res = {"model": {"geometry": "Llama-3-8B (n_embd 4096, n_ff 14336, 32/8 heads)", "layers": a.layers, "vocab": a.vocab, "type": a.type,
                     "weights": "random-init, quantized by ggml_quantize_chunk"},
           "n_prompt": a.prompt, "n_gen": a.gen, "warmup_tokens": a.warmup, "flags": "ngl 99, offload_kqv, flash attention, batch 512", "runs": {}}
    toks = {}
    for name, lib, env in configs:
        r = run(common + ["--backend", lib], env=env)
        if "error" in r:
            res["runs"][name] = r
            continue
        toks[name] = r["tokens"]
        res["runs"][name] = {"pp_tok_s": a.prompt / r["prompt_s"], "tg_tok_s": (a.gen - 1) / r["gen_s"], "ms_per_token": 1e3 * r["gen_s"] / (a.gen - 1),
                             "prompt_s": r["prompt_s"], "gen_s": r["gen_s"], "load_s": r["load_s"], "graph_splits": r["graph_splits"], "env": env}
        if "backend_timing" in r:
            res["runs"][name]["backend_timing"] = r["backend_timing"]
    if "b200" in toks and "reference_cuda" in toks:
        b, c = res["runs"]["b200"], res["runs"]["reference_cuda"]
        res["tg_speedup_vs_reference_cuda"] = b["tg_tok_s"] / c["tg_tok_s"]
        res["pp_speedup_vs_reference_cuda"] = b["pp_tok_s"] / c["pp_tok_s"]
    out = json.dumps(res, indent=1)
    print(out)
    if a.out:
        Path(a.out).write_text(out)
    for f in tmp.iterdir():
        f.unlink()
    tmp.rmdir()
    return 0


if __name__ == "__main__":
    sys.exit(main())
