#!/usr/bin/env python3
"""moe_bench.py — MUL_MAT_ID on Mixtral-8x7B expert shapes (BASELINE config 5): 8 experts, top-2 routing, q4_0, batch 1..64 tokens.
Prints per batch size the device time of one gate-sized (14336 x 4096) and one down-sized (4096 x 14336) MUL_MAT_ID, the
distinct weight bytes the routing touches, and the resulting effective GB/s. CUDA events on the launching stream."""
import argparse
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="q4_0")
    ap.add_argument("--batches", default="1,2,4,8,16,32,64")
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    import torch
    import b200_pkg
    pkg = b200_pkg.load()
    dev = torch.device("cuda", 0)
    t = bench.TYPES[args.type]
    slab = torch.from_numpy(np.load(ROOT / "tests" / "golden" / "weight_slabs.npz")[args.type]).to(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    n_as, n_used = 8, 2
    stream = torch.cuda.Stream(device=dev)
    rows = []
    with torch.cuda.stream(stream):
        ctx = pkg.Context(0)
        for (M, K) in ((14336, 4096), (4096, 14336)):
            idx = torch.randint(0, slab.shape[0], (n_as * M, K // 256), generator=g, device=dev)
            W = slab[idx].reshape(n_as, M, -1).contiguous()
            for nt in [int(b) for b in args.batches.split(",")]:
                B = (torch.rand((nt, 1, K), generator=g, device=dev) * 2 - 1).contiguous()
                ids = torch.stack([torch.randperm(n_as, generator=g, device=dev)[:n_used] for _ in range(nt)]).to(torch.int32).contiguous()
                out = ctx.mul_mat_id(t, W, B, ids)
                stream.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(args.reps):
                    ctx.mul_mat_id(t, W, B, ids, out=out)
                e1.record(stream)
                e1.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / args.reps
                distinct = len(torch.unique(ids))
                wbytes = distinct * M * (K // 256) * bench.SEG_BYTES[args.type]
                rows.append(dict(M=M, K=K, tokens=nt, pairs=nt * n_used, experts_touched=distinct, us=us,
                                 distinct_weight_GBs=wbytes / us * 1e-3, pair_weight_GBs=nt * n_used * M * (K // 256) * bench.SEG_BYTES[args.type] / us * 1e-3))
                print(f"{args.type} {M}x{K} tokens={nt:3d} pairs={nt * n_used:4d} experts={distinct}: {us:9.1f} us  "
                      f"{rows[-1]['distinct_weight_GBs']:8.1f} GB/s of distinct weights ({rows[-1]['pair_weight_GBs']:8.1f} GB/s per-pair)")
    print(json.dumps(rows))


if __name__ == "__main__":
    main()
