"""Greedy-decode token-stream parity through the REAL llama.cpp graph and loaders (north_star: "identical
greedy-decode token streams on a random-init model").

oracle/llama_harness.cpp (test infrastructure, linked to the reference's own libllama compiled into oracle/_ref) writes a
random-init Llama-shaped GGUF with the reference's gguf API - matmul weights quantized by ggml_quantize_chunk, no_vocab tokenizer -
and runs llama_decode + argmax over it. The stream from ggml-cpu alone (strict x64 build: quantize_row_*_ref activations, generic
vec_dots) is compared with the stream from `libggml-b200.so -ngl 99` for every weight type of the north star and every loader of
the fork: llama_model_load_from_file, _from_buffer, _from_split_futures (src/llama.cpp:313, 328, 361). The prompt is 64 tokens, so
the prefill kernels are on the path, then >= 256 tokens are generated one at a time through the decode mat-vec kernels.
"""
import json
import os
import subprocess
from pathlib import Path

import pytest

from oracle.pyoracle import REF_DIR

HARNESS = REF_DIR / "llama_harness"
CPU_LIB = REF_DIR / "libggml-cpu-x64.so"
ROOT = Path(__file__).resolve().parent.parent


def _need_harness():
    if not HARNESS.exists() or not (REF_DIR / "libllama.so").exists():
        pytest.skip("oracle/_ref/llama_harness not built (make -C oracle llama, where /root/reference exists)")


def _run(args, timeout=1200, env=None):
    p = subprocess.run([str(HARNESS)] + [str(a) for a in args], capture_output=True, text=True, timeout=timeout, env=env)
    assert p.returncode == 0, f"llama_harness {' '.join(map(str, args))}\n{p.stdout[-2000:]}\n{p.stderr[-4000:]}"
    return json.loads(p.stdout.strip().splitlines()[-1])


def _gen(prefix, tname, **shape):
    args = ["gen", "--type", tname, "--out", prefix]
    for k, v in shape.items():
        args += [f"--{k.replace('_', '-')}", v]
    return _run(args)


def _decode(prefix, loader, backends, ngl=0, prompt=64, gen=256, env=None, extra=()):
    args = ["run", "--model", prefix, "--loader", loader, "--ngl", ngl, "--prompt", prompt, "--gen", gen]
    for b in backends:
        args += ["--backend", b]
    return _run(args + list(extra), env=env)


def _first_divergence(a, b):
    for i, (x, y) in enumerate(zip(a["tokens"], b["tokens"])):
        if x != y:
            return i
    return None if len(a["tokens"]) == len(b["tokens"]) else min(len(a["tokens"]), len(b["tokens"]))


def test_loaders_agree_on_cpu(tmp_path):
    """the harness itself, without a GPU: a small model through the three loaders on ggml-cpu gives one token stream"""
    _need_harness()
    prefix = tmp_path / "tiny"
    _gen(prefix, "q4_K", layers=2, embd=256, ff=512, heads=4, kv_heads=2, vocab=512, splits=2)
    runs = {ld: _decode(prefix, ld, [CPU_LIB], prompt=16, gen=24, extra=["--splits", 2, "--threads", 2]) for ld in ("file", "buffer", "futures")}
    assert len(runs["file"]["tokens"]) == 24
    assert runs["file"]["tokens"] == runs["buffer"]["tokens"] == runs["futures"]["tokens"]
    assert len(set(runs["file"]["tokens"])) > 4   # a degenerate model would repeat one token


@pytest.mark.gpu
@pytest.mark.parametrize("tname", ["q4_K", "q4_0", "q8_0"])
def test_greedy_stream_matches_ggml_cpu(pkg, tmp_path, tname):
    """4 layers of Llama-3-8B geometry (n_embd 4096, n_ff 14336, GQA 32/8): 64-token prompt + 256 generated tokens, B200 backend
    against ggml-cpu, for the file / memory-buffer / split-future loaders"""
    _need_harness()
    if os.environ.get("B200_GREEDY_TYPES") and tname not in os.environ["B200_GREEDY_TYPES"].split(","):
        pytest.skip("not in B200_GREEDY_TYPES")
    if not pkg.BACKEND_PATH.exists():
        pytest.fail("libggml-b200.so missing: build it with __graft_entry__.build()")
    prefix = tmp_path / f"parity_{tname}"
    _gen(prefix, tname)
    n_gen = int(os.environ.get("B200_GREEDY_TOKENS", "256"))
    cpu = _decode(prefix, "file", [CPU_LIB], gen=n_gen)
    assert len(cpu["tokens"]) == n_gen and len(set(cpu["tokens"])) > 16
    report = {"type": tname, "n_prompt": 64, "n_gen": n_gen, "cpu": {k: cpu[k] for k in ("load_s", "prompt_s", "gen_s")},
              "min_margin_cpu": min(cpu["margins"]), "loaders": {}}
    failures = []
    for loader in ("file", "buffer", "futures"):
        got = _decode(prefix, loader, [CPU_LIB, pkg.BACKEND_PATH], ngl=99, gen=n_gen)
        d = _first_divergence(cpu, got)
        report["loaders"][loader] = {"identical": d is None, "first_divergence": d, "load_s": got["load_s"], "prompt_s": got["prompt_s"],
                                     "gen_s": got["gen_s"], "max_margin_delta": max(abs(a - b) for a, b in zip(cpu["margins"], got["margins"]))}
        if d is not None:
            failures.append(f"{tname}/{loader}: streams diverge at generated token {d}: cpu {cpu['tokens'][d]} vs b200 {got['tokens'][d]}; "
                            f"top-2 logit margin there: cpu {cpu['margins'][d]:.3g}, b200 {got['margins'][d]:.3g}")
    out = ROOT / "gpurun_out"
    if out.exists():
        (out / f"greedy_{os.environ.get('B200_GREEDY_TAG', 'default')}_{tname}.json").write_text(json.dumps(report, indent=1))
    assert not failures, "\n".join(failures)
