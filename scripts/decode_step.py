#!/usr/bin/env python3
"""decode_step.py — one WHOLE Llama-3-8B token step through the C-ABI (what libggml-b200.so launches for a decode cgraph once its
planner has fused it: per layer [RMS_NORM+MUL+q/k/v+ROPE q,k+KV-cache rows], [FLASH_ATTN_EXT], [attn_output+residual],
[RMS_NORM+MUL+gate/up+SWIGLU], [ffn_down+residual]; then the output norm and lm_head), replayed as a CUDA graph. Prints ms per token
and, with --skip, the same step without one class of launches: the marginal in-graph cost of each kernel (a dependent chain cannot
be timed kernel by kernel any other way without a profiler serialising it). Weights: tests/golden slabs gathered at random, like bench.py.

    python scripts/decode_step.py --n-kv 512 --skip none,prep,fa,norm,qkv,o,gateup,down,head
"""
import argparse
import json
import math
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="q4_K")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--vocab", type=int, default=32000)
    ap.add_argument("--n-kv", type=int, default=512)
    ap.add_argument("--n-ctx", type=int, default=4096)
    ap.add_argument("--skip", default="none")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--prefetch-kb", type=int, default=64)
    ap.add_argument("--one-launch-attn", action="store_true", help="b200_attn_decode instead of attn_prep + flash_attn_ext (measured slower)")
    ap.add_argument("--no-early-kv", action="store_true", help="plain b200_flash_attn_ext instead of b200_flash_attn_ext_decode")
    ap.add_argument("--no-qkv-rope", action="store_true", help="q/k/v and attn_prep as two launches instead of b200_qkv_rope's one")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import torch
    import b200_pkg
    pkg = b200_pkg.load()
    dev = torch.device("cuda", 0)
    t = bench.TYPES[a.type]
    E, FF, NH, NKV, D = 4096, 14336, 32, 8, 128
    slab = torch.from_numpy(bench.load_slab(a.type)).to(dev)
    g = torch.Generator(device=dev).manual_seed(1)

    def weight(M, K):
        idx = torch.randint(0, slab.shape[0], (M, K // 256), generator=g, device=dev)
        return slab[idx].reshape(M, -1).contiguous()
    L = []
    for _ in range(a.layers):
        L.append(dict(wq=weight(E, E), wk=weight(NKV * D, E), wv=weight(NKV * D, E), wo=weight(E, E), wg=weight(FF, E), wu=weight(FF, E), wd=weight(E, FF),
                      an=torch.rand(E, device=dev, generator=g) + 0.5, fn=torch.rand(E, device=dev, generator=g) + 0.5,
                      kc=(torch.randn((a.n_ctx, NKV * D), device=dev, generator=g) * 0.3).half(), vc=(torch.randn((a.n_ctx, NKV * D), device=dev, generator=g) * 0.3).half()))
    w_head, on = weight(a.vocab, E), torch.rand(E, device=dev, generator=g) + 0.5
    x0 = torch.randn((1, E), device=dev, generator=g)
    pos = torch.tensor([a.n_kv - 1], device=dev, dtype=torch.int32)
    idx = torch.tensor([a.n_kv - 1], device=dev, dtype=torch.int64)
    mask = torch.zeros((1, 1, 64, a.n_kv), device=dev, dtype=torch.float16)
    scale = 1.0 / math.sqrt(D)
    results = {}
    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        ctx = pkg.Context(0)
        ctx.prefetch(a.prefetch_kb)
        for skip in a.skip.split(","):
            def step():
                x = x0
                for l in L:
                    folded = False
                    if skip not in ("qkv", "prep", "attn", "norm") and not a.no_qkv_rope and not a.one_launch_attn:
                        # norm + q/k/v + ROPE + KV-cache rows: ONE launch
                        q4, _ = ctx.qkv_rope(t, [l["wq"], l["wk"], l["wv"]], x, pos, l["kc"], idx, l["vc"], idx, NH, NKV, D, 0, norm_w=l["an"], norm_eps=1e-5,
                                             freq_base=500000.0, n_ctx_orig=8192)
                        folded = True
                    elif skip != "qkv":
                        if skip == "norm":
                            q, k, v = ctx.mul_mat_shared(t, [l["wq"], l["wk"], l["wv"]], x)
                        else:
                            q, k, v = ctx.mul_mat_shared(t, [l["wq"], l["wk"], l["wv"]], x, norm_w=l["an"], norm_eps=1e-5)
                    else:
                        q, k, v = x0[:, :E], x0[:, :NKV * D], x0[:, NKV * D:2 * NKV * D]
                    if not folded:
                        q4, k4 = q.reshape(1, 1, NH, D), k.reshape(1, 1, NKV, D)
                    if folded:
                        kview = l["kc"][:a.n_kv].view(1, a.n_kv, NKV, D).permute(0, 2, 1, 3)
                        vview = l["vc"][:a.n_kv].view(1, a.n_kv, NKV, D).permute(0, 2, 1, 3)
                        att = ctx.flash_attn_ext(q4.permute(0, 2, 1, 3), kview, vview, mask, None, scale, written_idx=None if a.no_early_kv else idx).reshape(1, E) if skip != "fa" else q4.reshape(1, E)
                    elif skip == "attn":
                        att = q4.reshape(1, E)
                    elif skip not in ("prep", "fa") and a.one_launch_attn:   # ROPE + KV store + attention: one launch
                        att, _, _ = ctx.attn_decode(q4, k4, pos, l["kc"], idx, v.reshape(1, NKV * D), l["vc"], idx, a.n_kv, mask, D, 0, scale=scale,
                                                    freq_base=500000.0, n_ctx_orig=8192)
                        att = att.reshape(1, E)
                    else:
                        if skip != "prep":
                            q4, _ = ctx.attn_prep(q4, k4, pos, l["kc"], idx, v.reshape(1, NKV * D), l["vc"], idx, D, 0, freq_base=500000.0, n_ctx_orig=8192)
                        if skip != "fa":
                            kview = l["kc"][:a.n_kv].view(1, a.n_kv, NKV, D).permute(0, 2, 1, 3)
                            vview = l["vc"][:a.n_kv].view(1, a.n_kv, NKV, D).permute(0, 2, 1, 3)
                            att = ctx.flash_attn_ext(q4.permute(0, 2, 1, 3), kview, vview, mask, None, scale).reshape(1, E)
                        else:
                            att = q4.reshape(1, E)
                    x2 = ctx.mul_mat_fused(t, l["wo"], None, att, up_bias=x.reshape(-1)) if skip != "o" else x
                    if skip != "gateup":
                        if skip == "norm":
                            h = ctx.mul_mat_fused(t, l["wu"], l["wg"], x2, glu_op=pkg.GLU_SWIGLU)
                        else:
                            h = ctx.mul_mat_fused(t, l["wu"], l["wg"], x2, glu_op=pkg.GLU_SWIGLU, norm_w=l["fn"], norm_eps=1e-5)
                    else:
                        h = hbuf
                    x = ctx.mul_mat_fused(t, l["wd"], None, h, up_bias=x2.reshape(-1)) if skip != "down" else x2
                if skip != "head":
                    xn = ctx.rms_norm_mul(x.reshape(1, 1, 1, E), on, 1e-5).reshape(1, E)
                    return ctx.mul_mat(t, w_head, xn)
                return x
            hbuf = torch.randn((1, FF), device=dev, generator=g) * 0.1
            step()
            l0 = ctx.launches
            step()
            n_launch = ctx.launches - l0
            stream.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=stream):
                step()
            for _ in range(5):
                gr.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(a.steps):
                gr.replay()
            e1.record(stream)
            stream.synchronize()
            ms = e0.elapsed_time(e1) / a.steps
            results[skip] = {"ms_per_token": ms, "launches": n_launch}
            print(f"skip={skip:8s} {ms:8.4f} ms/token  {n_launch} launches", flush=True)
    if "none" in results:
        base = results["none"]["ms_per_token"]
        for k, v in results.items():
            if k != "none":
                v["marginal_us_per_layer"] = (base - v["ms_per_token"]) * 1e3 / (1 if k == "head" else a.layers)
                print(f"  {k:8s} costs {v['marginal_us_per_layer']:.2f} us per {'token' if k == 'head' else 'layer'} in the graph")
    out = {"type": a.type, "layers": a.layers, "vocab": a.vocab, "n_kv": a.n_kv, "results": results}
    if a.out:
        Path(a.out).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
