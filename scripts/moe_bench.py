#!/usr/bin/env python3
"""moe_bench.py — BASELINE configs 5 and 1, one JSON line each.

config 5: MUL_MAT_ID on Mixtral-8x7B expert shapes: 8 experts, top-2 routing, q4_0, batch 1..64 tokens. A step = the expert
  matmuls of ALL 32 layers (gate, up: 14336 x 4096 x 8 experts; down: 4096 x 14336 x 8 experts: 25.4 GB of expert weights resident,
  far beyond the 126 MB L2), one MUL_MAT_ID launch group per tensor as in the llama graph (src/llama-graph.cpp:1038-1096), replayed as
  a CUDA graph, CUDA events on the launching stream. Reported: ms per step, GB/s of the DISTINCT weights the routing touches
  (the HBM-roofline figure: algorithmic bytes = distinct experts x M x K x bpw + activations), tokens/s of the MoE stack.
config 1: one MUL_MAT q4_0 4096 x 4096, n = 1 (9 469 952 algorithmic bytes), two ways: the SAME matrix every launch (what
  test-backend-ops perf does: L2-resident on this part) and 128 distinct matrices in rotation (1.2 GB: from HBM)."""
import argparse
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def timed(torch, stream, fn, reps, graph=True):
    fn()
    stream.synchronize()
    if not graph:   # prefill-sized MoE batches read the expert bounds back on the host: not capturable, timed as eager launches
        fn()
        stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        e1.synchronize()
        return e0.elapsed_time(e1) / reps
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        fn()
    for _ in range(3):
        g.replay()
    stream.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        g.replay()
    e1.record(stream)
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="q4_0")
    ap.add_argument("--batches", default="1,2,4,8,16,32,64,512")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import b200_pkg
    pkg = b200_pkg.load()
    dev = torch.device("cuda", 0)
    t = bench.TYPES[args.type]
    slab = torch.from_numpy(bench.load_slab(args.type)).to(dev)
    seg = bench.SEG_BYTES[args.type]
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    n_as, n_used, E, F = 8, 2, 4096, 14336
    stream = torch.cuda.Stream(device=dev)
    peak, _ = bench.hbm_peak()
    lines = []
    with torch.cuda.stream(stream):
        ctx = pkg.Context(0)
        ctx.prefetch(0)

        def experts(M, K):
            idx = torch.randint(0, slab.shape[0], (n_as * M, K // 256), generator=g, device=dev)
            return slab[idx].reshape(n_as, M, -1).contiguous()
        layers = [dict(gate=experts(F, E), up=experts(F, E), down=experts(E, F)) for _ in range(args.layers)]
        for nt in [int(b) for b in args.batches.split(",")]:
            x_in = (torch.rand((nt, 1, E), generator=g, device=dev) * 2 - 1).contiguous()
            x_mid = (torch.rand((nt, n_used, F), generator=g, device=dev) * 2 - 1).contiguous()
            ids = [torch.stack([torch.randperm(n_as, generator=g, device=dev)[:n_used] for _ in range(nt)]).to(torch.int32).contiguous()
                   for _ in range(args.layers)]
            o_g = torch.empty((nt, n_used, F), dtype=torch.float32, device=dev)
            o_u = torch.empty_like(o_g)
            o_d = torch.empty((nt, n_used, E), dtype=torch.float32, device=dev)

            def step():
                for l, L in enumerate(layers):
                    ctx.mul_mat_id(t, L["gate"], x_in, ids[l], out=o_g)
                    ctx.mul_mat_id(t, L["up"], x_in, ids[l], out=o_u)
                    ctx.mul_mat_id(t, L["down"], x_mid, ids[l], out=o_d)
            l0 = ctx.launches
            step()
            launches = ctx.launches - l0
            tiled = nt * n_used >= 48 * n_as
            ms = timed(torch, stream, step, args.reps, graph=not tiled)
            distinct = sum(len(torch.unique(i)) for i in ids)
            wbytes = distinct * 3 * F * (E // 256) * seg
            abytes = args.layers * (2 * (nt * E * 4 + nt * n_used * F * 4) + nt * n_used * F * 4 + nt * n_used * E * 4)
            gbs = (wbytes + abytes) / (ms * 1e-3) / 1e9
            line = {"metric": f"{args.type} MUL_MAT_ID HBM GB/s of distinct expert weights, Mixtral-8x7B expert shapes, top-2 of 8", "value": gbs,
                    "unit": "GB/s", "ms_per_step": ms, "tokens": nt, "pairs_per_layer": nt * n_used, "tok_s": nt / (ms * 1e-3),
                    "experts_touched_avg": distinct / args.layers, "launches_per_step": launches,
                    "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "frac": gbs / peak, "unit": "GB/s"},
                    "config": {"workload": f"Mixtral-8x7B MoE MUL_MAT_ID stack, {args.layers} layers x (gate, up, down) x 8 experts, {args.type}, "
                                           f"{nt} tokens, top-2", "launch": "eager launches (per-expert tile GEMMs: the expert bounds are read back on the host)" if tiled else "CUDA graph replay", "resident_expert_bytes": args.layers * 3 * n_as * F * (E // 256) * seg}}
            lines.append(line)
            print(json.dumps(line), flush=True)
        # ---- config 1
        M = K = 4096
        one = slab[torch.randint(0, slab.shape[0], (M, K // 256), generator=g, device=dev)].reshape(M, -1).contiguous()
        many = [slab[torch.randint(0, slab.shape[0], (M, K // 256), generator=g, device=dev)].reshape(M, -1).contiguous() for _ in range(128)]
        x = (torch.rand((1, K), generator=g, device=dev) * 2 - 1).contiguous()
        out = torch.empty((1, M), dtype=torch.float32, device=dev)
        nbytes = bench.algorithmic_bytes(args.type, M, K, 1)
        ms_same = timed(torch, stream, lambda: [ctx.mul_mat(t, one, x, out=out) for _ in range(128)], args.reps) / 128
        ms_rot = timed(torch, stream, lambda: [ctx.mul_mat(t, w, x, out=out) for w in many], args.reps) / 128
        line = {"metric": f"{args.type} MUL_MAT 4096x4096 n=1 (BASELINE config 1)", "unit": "GB/s", "algorithmic_bytes": nbytes,
                "same_matrix_every_launch": {"us": ms_same * 1e3, "value": nbytes / (ms_same * 1e-3) / 1e9, "note": "L2-resident, the reference's test-backend-ops perf pattern"},
                "128_distinct_matrices": {"us": ms_rot * 1e3, "value": nbytes / (ms_rot * 1e-3) / 1e9, "frac_of_hbm_peak": nbytes / (ms_rot * 1e-3) / 1e9 / peak,
                                          "note": "1.2 GB in rotation: every launch streams from HBM; launches are dependent (stream order, PDL)"}}
        lines.append(line)
        print(json.dumps(line), flush=True)
    if args.out:
        Path(args.out).write_text("\n".join(json.dumps(l) for l in lines) + "\n")


if __name__ == "__main__":
    main()
