#!/usr/bin/env python3
"""layer_timeline.py — device-side timeline (b200_ctx_trace, %globaltimer stamps written by the kernels themselves, no profiler) of the
five launches of a decode layer inside a replayed CUDA graph: [norm+q/k/v+rope+KV store] [FLASH_ATTN_EXT] [attn_output+residual]
[norm+gate/up+SWIGLU] [down+residual]. Per launch, over its CTAs, in us from the first stamp of the printed window:
start = CTA resident; dep = past griddepcontrol.wait (the previous kernel has completed); mid = mat-vec: activations quantized /
attention: first round's rows used; work = mat-vec: first weight stage landed / attention: rounds done; done = last stamp."""
import argparse
import math
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="q4_K")
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--n-kv", type=int, default=512)
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
                      kc=(torch.randn((a.n_kv + 64, NKV * D), device=dev, generator=g) * 0.3).half(),
                      vc=(torch.randn((a.n_kv + 64, NKV * D), device=dev, generator=g) * 0.3).half()))
    x0 = torch.randn((1, E), device=dev, generator=g)
    pos = torch.tensor([a.n_kv - 1], device=dev, dtype=torch.int32)
    idx = torch.tensor([a.n_kv - 1], device=dev, dtype=torch.int64)
    mask = torch.zeros((1, 1, 64, a.n_kv), device=dev, dtype=torch.float16)
    scale = 1.0 / math.sqrt(D)
    names = ["norm+qkv+rope", "flash_attn", "o+res", "norm+gate/up+glu", "down+res"]
    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        ctx = pkg.Context(0)
        ctx.prefetch(64)

        def step():
            x = x0
            for l in L:
                q4, _ = ctx.qkv_rope(t, [l["wq"], l["wk"], l["wv"]], x, pos, l["kc"], idx, l["vc"], idx, NH, NKV, D, 0, norm_w=l["an"], norm_eps=1e-5,
                                     freq_base=500000.0, n_ctx_orig=8192)
                kview = l["kc"][:a.n_kv].view(1, a.n_kv, NKV, D).permute(0, 2, 1, 3)
                vview = l["vc"][:a.n_kv].view(1, a.n_kv, NKV, D).permute(0, 2, 1, 3)
                att = ctx.flash_attn_ext(q4.permute(0, 2, 1, 3), kview, vview, mask, None, scale, written_idx=idx).reshape(1, E)
                x2 = ctx.mul_mat_fused(t, l["wo"], None, att, up_bias=x.reshape(-1))
                h = ctx.mul_mat_fused(t, l["wu"], l["wg"], x2, glu_op=pkg.GLU_SWIGLU, norm_w=l["fn"], norm_eps=1e-5)
                x = ctx.mul_mat_fused(t, l["wd"], None, h, up_bias=x2.reshape(-1))
            return x
        step()
        l0 = ctx.launches
        step()
        n = ctx.launches - l0
        stream.synchronize()
        ctx.trace(n)          # zeroed; nothing runs eagerly from here on, so every slot is written by exactly one launch of the graph
        c0 = ctx.launches     # launch i of the step stamps slot (c0 + i) % n
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=stream):
            step()
        for _ in range(5):
            gr.replay()
        stream.synchronize()
        tr = ctx.trace_read().astype(np.int64)   # [n, 148, 8]
    order = [(c0 + i) % n for i in range(n)]
    order = [s for s in order if (tr[s, :, 0] > 0).any()]
    t0 = min(tr[s, :, 0][tr[s, :, 0] > 0].min() for s in order)
    name_of = {(c0 + i) % n: names[i % 5] for i in range(n)}
    lines = [f"{'launch':<18}{'ctas':>5} | {'start':>13} | {'dep':>13} | {'x loaded':>13} | {'mid':>13} | {'work':>13} | {'done':>13} | since previous done"]
    prev = None
    for k, s in enumerate(order):
        live = tr[s, :, 0] > 0
        cols = []
        for i in (0, 3, 7, 4, 5, 6):
            v = tr[s, live, i]
            v = v[v > 0]
            cols.append(f"{(v.min() - t0) / 1e3:6.2f}-{(v.max() - t0) / 1e3:6.2f}" if v.size else "      -      ")
        end = tr[s, live, :].max()
        lines.append(f"{name_of[s]:<18}{int(live.sum()):>5} | " + " | ".join(cols) + f" | {(end - prev) / 1e3 if prev else 0:6.2f} us")
        prev = end
    txt = "\n".join(lines)
    print(txt)
    if a.out:
        Path(a.out).write_text(txt + "\n")


if __name__ == "__main__":
    main()
