#!/usr/bin/env python3
"""mmf_timeline.py — device-side timeline of the prefill GEMM launches of one Llama-3-8B layer (b200_ctx_trace, no profiler):
per launch the spread over CTAs of CTA start / set-up done + dependency resolved / first TMA issued / first stage landed /
CTA done (every role through its span) / last MMA of the first work item issued / its accumulator complete / its epilogue done, in us."""
import argparse
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="q4_K")
    ap.add_argument("--tokens", type=int, default=512)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import b200_pkg
    pkg = b200_pkg.load()
    device = torch.device("cuda", 0)
    m = dict(bench.MODELS["8b"])
    m["n_layer"] = 1
    t = bench.TYPES[args.type]
    stream = torch.cuda.Stream(device=device)
    with torch.cuda.stream(stream):
        ctx = pkg.Context(0)
        ctx.weight_cache(True)
        ops = bench.build_stack(torch, m, args.type, args.tokens, 1, 0, device)

        groups = bench.group_ops(ops)

        def step():   # like bench.py: MUL_MATs that share src1 go through b200_mul_mat_shared_src1 (one grouped launch)
            for grp in groups:
                if len(grp) == 1:
                    ctx.mul_mat(t, grp[0]["W"], grp[0]["x"], out=grp[0]["out"])
                else:
                    ctx.mul_mat_shared(t, [o["W"] for o in grp], grp[0]["x"], outs=[o["out"] for o in grp])

        step()
        l0 = ctx.launches
        step()
        n = ctx.launches - l0
        stream.synchronize()
        ctx.trace(n)
        step()
        run = step
        if not args.no_graph:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                step()
            run = g.replay
        for _ in range(3):
            run()
        stream.synchronize()
        tr = ctx.trace_read().astype(np.int64)          # [n, 148, 8]
    used = [s for s in range(n) if (tr[s, :, 0] > 0).any()]
    starts = np.array([tr[s, :, 0][tr[s, :, 0] > 0].min() for s in used])
    order = [used[i] for i in np.argsort(starts)]
    t0 = starts.min()
    lines = [f"{'launch':<14}{'ctas':>5} | {'start':>13} | {'dep ok':>13} | {'parts 0 in':>13} | {'stage 0':>13} | {'cta done':>13} | {'last mma':>13} | {'drained':>13} | {'epi done':>13}"]
    for k, s in enumerate(order):
        live = tr[s, :, 0] > 0
        cols = []
        for i in range(8):
            v = tr[s, live, i]
            v = v[v > 0]
            cols.append(f"{(v.min() - t0) / 1e3:6.2f}-{(v.max() - t0) / 1e3:6.2f}" if v.size else "      -      ")
        name = "+".join(o["name"] for o in groups[k % len(groups)])
        lines.append(f"{name[:13]:<14}{int(live.sum()):>5} | " + " | ".join(cols))
    txt = "\n".join(lines)
    print(txt)
    if args.out:
        Path(args.out).write_text(txt + "\n")


if __name__ == "__main__":
    main()
