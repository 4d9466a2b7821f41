#!/usr/bin/env python3
"""mv_timeline.py — device-side timeline of a decode step's mat-vec launches (b200_ctx_trace), no profiler attached.
Prints, per launch of one CUDA-graph replay, the spread over CTAs of: CTA start, first/last weight copy issued, consumers
past the stream dependency, activations quantized, first stage landed, warp 0 done — in us relative to the step's start."""
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
    ap.add_argument("--layers", type=int, default=3)
    ap.add_argument("--workload", default="8b")
    ap.add_argument("--prefetch-kb", type=int, default=0)
    ap.add_argument("--ncols", type=int, default=1)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--split-selftest", action="store_true", help="row-split epilogue against this same GPU (peer flag pre-armed): local cost of the exchange")
    ap.add_argument("--deferred", action="store_true")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import b200_pkg
    pkg = b200_pkg.load()
    device = torch.device("cuda", 0)
    m = dict(bench.MODELS[args.workload])
    m["n_layer"] = args.layers
    t = bench.TYPES[args.type]
    stream = torch.cuda.Stream(device=device)
    lines = []
    with torch.cuda.stream(stream):
        ctx = pkg.Context(0)
        if args.prefetch_kb > 0:
            ctx.prefetch(args.prefetch_kb)
        if args.split_selftest:
            sync = torch.zeros(64, dtype=torch.int32, device=device)
            sync[1] = 0x7FFFFFF0
            torch.cuda.synchronize()
            dummy = torch.zeros(16, dtype=torch.float32, device=device)
            ctx.row_split(0, [dummy.data_ptr(), dummy.data_ptr()], [sync.data_ptr(), sync.data_ptr() + 128])
            ctx.row_split_deferred(args.deferred)
        ops = bench.build_stack(torch, m, args.type, args.ncols, 1, 0, device)
        groups = bench.group_ops(ops)

        def step():
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
        for _ in range(5):
            run()
        stream.synchronize()
        tr = ctx.trace_read().astype(np.int64)          # [n, 148, 8]
    names = ["+".join(o["name"] for o in grp) for grp in groups]
    starts = np.array([tr[s, :, 0][tr[s, :, 0] > 0].min() for s in range(n)])
    order = np.argsort(starts)
    t0 = starts.min()
    hdr = f"{'launch':<28}{'ctas':>5} | {'start':>13} | {'pre copy0':>13} | {'post copy0':>13} | {'dep ok':>13} | {'x loaded':>13} | {'quantized':>13} | {'stage 0':>13} | {'w0 done':>13} | MB    GB/s(span)"
    lines.append(hdr)
    prev_end = None
    # slots rotate with the launch counter: the i-th launch in time order is group i of the step
    for k, s in enumerate(order):
        live = tr[s, :, 0] > 0
        cols = []
        for i in (0, 1, 2, 3, 7, 4, 5, 6):
            v = tr[s, live, i]
            v = v[v > 0]
            cols.append(f"{(v.min() - t0) / 1e3:6.2f}-{(v.max() - t0) / 1e3:6.2f}" if v.size else "      -      ")
        grp = groups[k % len(groups)]
        mb = sum(bench.algorithmic_bytes(args.type, o["Ml"], o["K"], args.ncols) for o in grp) / 1e6
        end = tr[s, live, 6].max()
        span = (end - (prev_end if prev_end is not None else tr[s, live, 0].min())) / 1e3
        prev_end = end
        lines.append(f"{names[k % len(names)][:27]:<28}{int(live.sum()):>5} | " + " | ".join(cols) + f" | {mb:5.1f} {mb / span * 1e-3 if span > 0 else 0:6.2f} TB/s ({span:5.2f} us)")
    txt = "\n".join(lines)
    print(txt)
    if args.out:
        Path(args.out).write_text(txt + "\n")


if __name__ == "__main__":
    main()
